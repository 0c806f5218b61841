"""smoke(): one small pass of the hot path on cuda:0 through the C ABI, checked bit-exactly
against the CPU oracle. Used by __graft_entry__.smoke(); the oracle is imported here only as
the checker (it is test infrastructure and never on the product path)."""
from __future__ import annotations

from . import abi, synth
from .abi import Intersect, Row, Union, Xor, Difference
from .engine import Engine
from .executor import Executor


def smoke() -> None:
    import oracle  # checker only

    eng = Engine(0)
    orc = oracle.Oracle(workers=4)
    ex_g, ex_c = Executor(eng), Executor(orc)
    shards = [0, 1, 2]
    for ex in (ex_g, ex_c):
        ex.create_field("f")
        ex.create_field("m")
        ex.create_field("z")
        ex.create_field("v", type="int", min=-(1 << 15), max=(1 << 15) - 1)
    for s in shards:
        dense = synth.bernoulli(s, 0, 4, 0.10)
        mixed = synth.mixed(s, 0, 4)
        zipf = synth.zipf(s, n_rows=200, k=4)
        bsi = synth.bsi(s, bit_depth=16, vmin=-(1 << 15), vmax=(1 << 15) - 1, density=0.5)
        for ex in (ex_g, ex_c):
            ex.load_fragment("f", s, dense)
            ex.load_fragment("m", s, mixed)
            ex.load_fragment("z", s, zipf)
            ex.load_fragment("v", s, bsi)

    checks = [
        Intersect(Row("f", 1), Row("f", 2)),
        Union(Row("m", 0), Row("m", 1), Row("m", 2), Row("m", 3)),
        Difference(Row("m", 0), Row("m", 1), Row("f", 2), Row("m", 3)),
        Xor(Row("m", 0), Row("f", 1), Row("m", 2), Row("m", 3)),
        Intersect(Union(Row("m", 0), Row("f", 1)), Union(Row("m", 2), Row("f", 3))),
    ]
    for call in checks:
        g, c = ex_g.count(call, shards), ex_c.count(call, shards)
        assert g == c, f"Count({call}): gpu={g} oracle={c}"
    src = Intersect(Row("f", 1), Row("f", 2))
    g, c = ex_g.topn("z", shards, n=10, src=src), ex_c.topn("z", shards, n=10, src=src)
    assert g == c, f"TopN: gpu={g} oracle={c}"
    for name in ("sum", "min", "max"):
        g, c = getattr(ex_g, name)("v", shards), getattr(ex_c, name)("v", shards)
        assert g == c, f"{name}: gpu={g} oracle={c}"
    rc = ex_g.row_range("v", abi.BSI_GT, 1234)
    g, c = ex_g.count(rc, shards), ex_c.count(ex_c.row_range("v", abi.BSI_GT, 1234), shards)
    assert g == c, f"Count(Row(v > 1234)): gpu={g} oracle={c}"
    t = eng.last_timing()
    print(f"smoke ok: count/topn/bsi match the oracle bit-exactly; last call kernel {t.kernel_ms:.3f} ms, {t.launches} launches")
    eng.close()
