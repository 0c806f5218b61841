"""GPU test of pb200_count_pairs_noalloc, the entry point for hosts that may not allocate
device memory while serving (a host inside a PyTorch/RMM process whose pool owns the HBM, a
deployment whose whole HBM budget went to fragments, a query thread that must never block in
the driver's allocator). Its scratch is a fixed array inside the library, shared by every
call on the device, so the test calls it twice in a row with different inputs — the second
call must not see anything the first one left behind — and checks both answers against the
ordinary allocating entry point (pb200_count_batch) and, so that the two engine paths cannot
agree on a common mistake, against the CPU oracle holding the same fragments; integer
results, so the tolerance is exact equality, as everywhere else in this suite.

Named test_zz_* so pytest collects it last: if it fails under -x every other test has
already run. It builds its own engine and index, so it also passes alone in a fresh process.
"""
import random

import pytest

pytestmark = pytest.mark.gpu


def _expected(eng, fam, pairs, shards):
    from pilosa_b200 import abi

    progs = [abi.flatten(abi.Intersect(abi.Row("f", a), abi.Row("f", b)), lambda f, v: fam) for a, b in pairs]
    return [int(v) for v in eng.count_batch(progs, shards)[: len(pairs)]]


def test_count_pairs_noalloc_matches_allocating_entry_point_twice_in_a_row():
    import torch

    from pilosa_b200 import synth
    from pilosa_b200.engine import Engine, EngineError

    eng = Engine(0)
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    fam = eng.family_create()
    import oracle  # the checker; also used to splice two generated fragments into one file image

    orc = oracle.Oracle(workers=2)
    ofam = orc.family_create()
    shards = [0, 1, 2, 4]
    for s in shards:
        # rows 0-7: mixed array / bitmap / run containers (exercises the deferred-item list that
        # lives in the fixed scratch array); rows 8-11 dense bitmaps. One fragment per shard.
        a = oracle.Bitmap.deserialize(synth.mixed(s, 0, 8).bytes())
        b = oracle.Bitmap.deserialize(synth.bernoulli(s, 8, 4, 0.10).bytes())
        image = a.union(b).serialize()
        eng.frag_load(fam, s, image)
        orc.frag_load(ofam, s, image)
    rnd = random.Random(11)
    rows = list(range(12)) + [40]  # 40: a row that does not exist
    batches = [[(rnd.choice(rows), rnd.choice(rows)) for _ in range(n)] for n in (37, 64)]
    # the ordinary entry point first: it also refreshes the device-side fragment tables,
    # which the allocation-free path refuses to do itself
    expected = [_expected(eng, fam, pairs, shards + [3]) for pairs in batches]
    from pilosa_b200 import abi

    for pairs, exp in zip(batches, expected):  # the allocating entry point itself agrees with the oracle
        want = [orc.count(abi.flatten(abi.Intersect(abi.Row("f", a), abi.Row("f", b)), lambda f, v: ofam), shards + [3]) for a, b in pairs]
        assert exp == want

    with torch.cuda.stream(stream):
        d_shards = torch.tensor(shards + [3], dtype=torch.int64, device="cuda")  # shard 3 has no fragment
        for pairs, exp in zip(batches, expected):  # two calls in a row on the shared scratch array
            d_pairs = torch.tensor([x for p in pairs for x in p], dtype=torch.int64, device="cuda")
            d_counts = torch.full((len(pairs),), -1, dtype=torch.int64, device="cuda")  # garbage the call must overwrite
            stream.synchronize()
            eng.count_pairs_noalloc(fam, d_pairs.data_ptr(), len(pairs), d_shards.data_ptr(), d_shards.numel(), d_counts.data_ptr())
            stream.synchronize()
            assert d_counts.tolist() == exp
        # sizes beyond the fixed scratch array are rejected with a clear error, nothing is launched
        big = torch.zeros(2 * 4097, dtype=torch.int64, device="cuda")
        out = torch.zeros(4097, dtype=torch.int64, device="cuda")
        with pytest.raises(EngineError, match="PB200_NOALLOC_MAX_PAIRS"):
            eng.count_pairs_noalloc(fam, big.data_ptr(), 4097, d_shards.data_ptr(), 1, out.data_ptr())
        many = torch.zeros(70000, dtype=torch.int64, device="cuda")
        with pytest.raises(EngineError, match="PB200_NOALLOC_MAX_ITEMS"):
            eng.count_pairs_noalloc(fam, big.data_ptr(), 1, many.data_ptr(), 70000, out.data_ptr())
        stream.synchronize()
    eng.close()
