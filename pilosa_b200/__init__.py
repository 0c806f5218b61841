"""pilosa_b200 — B200 (sm_100a) execution engine for Pilosa's per-shard bitmap query hot path.

The product is the C-ABI shared library built from ``pilosa_b200/csrc`` (see
``include/pilosa_b200.h``); this package is the thin host-side mirror of the reference's
executor / fragment / Row call surface used by the tests, the bench and multi-GPU runs.
"""
from .abi import (  # noqa: F401
    Call, Difference, Empty, Intersect, Not, Row, Shift, Union, Xor, flatten,
)

__all__ = ["Call", "Row", "Intersect", "Union", "Difference", "Xor", "Not", "Shift", "Empty", "flatten"]
