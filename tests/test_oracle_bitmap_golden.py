"""Bitmap- and Row-level known answers of the reference (roaring/roaring_test.go, row_test.go),
restated literally, plus file-format and differential (naive sorted-set) checks of the oracle."""
import random

import oracle
from oracle import Bitmap, Row

SW = 1 << 20


def _both(a, b, n):
    assert a.intersection_count(b) == n and b.intersection_count(a) == n
    assert a.intersect(b).count() == n


# ---- roaring/roaring_test.go:980-1084 TestBitmap_IntersectionCount_* ----
def test_intersection_count_known_answers():
    _both(Bitmap(0, 1000001, 1000002, 1000003), Bitmap(0, 50000, 999998, 999999, 1000000, 1000001, 1000002), 3)  # ArrayArray
    b1 = Bitmap(0, 1, 2, 3, 4, 5, 1000000, 1000002, 1000003, 1000004, 1000005, 1000006)
    b1.optimize()
    assert {t for _, t, _ in b1.containers()} == {oracle.RUN}
    _both(Bitmap(0, 1000001, 1000002, 1000003), b1, 3)                                                           # ArrayRun
    b0 = Bitmap(3, 4, 5, 6, 7, 8, 1000001, 1000002, 1000003, 1000004)
    b0.optimize()
    _both(b0, b1, 6)                                                                                             # RunRun
    odd = Bitmap(*range(3, 1000007, 2))
    assert oracle.BITMAP in {t for _, t, _ in odd.containers()}
    _both(odd, b1, 4)                                                                                            # BitmapRun
    even = Bitmap(*range(0, 10001, 2))
    _both(Bitmap(1, 70, 200, 4097, 4098), even, 3)                                                               # ArrayBitmap
    x, y = Bitmap(*range(0, 10001, 2)), Bitmap(*range(1, 10002, 2))
    for v in (1000, 2000):
        x.add(v)
        y.add(v)
    _both(x, y, 2)                                                                                               # BitmapBitmap


# ---- roaring/roaring_test.go:1086-1111 TestBitmap_Shift ----
def test_bitmap_shift():
    mx = (1 << 64) - 1
    assert Bitmap(0, 1, 2, 3, 4, 5, 6, 7, 9, 10, 65536, mx).shift().slice() == [1, 2, 3, 4, 5, 6, 7, 8, 10, 11, 65537]
    assert Bitmap(65535, 131073).shift().slice() == [65536, 131074]
    assert Bitmap(65535, 131073, 65536 * 5 - 1, 65536 * 10, 65536 * 15 - 1).shift().slice() == \
        [65536, 131074, 65536 * 5, 65536 * 10 + 1, 65536 * 15]


# ---- row_test.go:26-127 ----
def test_row_merge_xor_union_difference():
    r1 = Row(1, 2, 3, SW + 1, 2 * SW)
    assert r1.merge(Row(3, 4, 5)).count() == 7
    assert Row().merge(Row(2, 66000, 70000, 70001, 70002, 70003, 70004)).count() == 7
    a, b = Row(0, 1, SW), Row(0, 2 * SW)
    assert a.xor(b).columns() == [1, SW, 2 * SW] and b.xor(a).columns() == [1, SW, 2 * SW]
    assert a.union(b).columns() == [0, 1, SW, 2 * SW] and b.union(a).columns() == [0, 1, SW, 2 * SW]
    assert a.difference(b).columns() == [1, SW]
    assert Row(0, 2 * SW).intersect(Row(1, SW)).count() == 0


# ---- file format: roaring_internal_test.go:1659-1751 (write/read per type), :2435-2454 (empties skipped) ----
def test_file_format_round_trip_and_ops_log():
    vals = [1, 5, 9, 70000, 70002] + list(range(200000, 200000 + 5000)) + list(range(400000, 460000, 2))
    b = Bitmap(*vals)
    data = b.serialize()
    assert data[:4] == (12348).to_bytes(2, "little") + b"\x00\x00"          # cookie: magic 12348, version 0, flags 0
    back = Bitmap.deserialize(data)
    assert back.slice() == sorted(vals) and back.serialize() == data
    types = {k: t for k, t, _ in back.containers()}
    assert types[0] == oracle.ARRAY and types[3] == oracle.RUN and types[6] == oracle.BITMAP
    # ops log appended to the snapshot is replayed on load (roaring.go:1627-1650)
    ops = oracle.encode_op(0, [11]) + oracle.encode_op(1, [5]) + oracle.encode_op(2, [500000, 500001]) + oracle.encode_op(3, [70000, 12345])
    after = Bitmap.deserialize(data + ops).slice()
    assert after == sorted((set(vals) | {11, 500000, 500001}) - {5, 70000})
    # corrupted checksum / bad magic are rejected with the reference's messages
    bad = bytearray(data + oracle.encode_op(0, [9]))
    bad[-1] ^= 0xFF
    for blob, msg in ((bytes(bad), "checksum"), (b"\x3c\x30\x01\x00" + b"\x00" * 12, "wrong roaring version"),
                      (b"\x00" * 16, "serialCookie"), (b"\x01\x02", "too small")):
        try:
            Bitmap.deserialize(blob)
            raise AssertionError("expected failure: " + msg)
        except oracle.OracleError as e:
            assert msg in str(e)
    assert Bitmap().serialize() == (12348).to_bytes(4, "little") + (0).to_bytes(4, "little")


# ---- differential check against the sorted-slice model (roaring/naive.go, fuzzer.go:96-163) ----
def test_bitmap_ops_vs_naive_sets():
    rnd = random.Random(5)

    def rand_bitmap():
        vals = set()
        for _ in range(rnd.randint(1, 4)):
            base = rnd.randrange(0, 6) << 16
            kind = rnd.random()
            if kind < 0.35:
                vals |= {base + rnd.randrange(65536) for _ in range(rnd.randint(1, 3000))}
            elif kind < 0.7:
                vals |= {base + v for v in range(65536) if rnd.random() < 0.3}
            else:
                s = rnd.randrange(60000)
                vals |= set(range(base + s, base + s + rnd.randint(1, 5000)))
        b = Bitmap(*sorted(vals))
        if rnd.random() < 0.7:
            b.optimize()
        return b, vals

    for _ in range(25):
        (a, sa), (b, sb) = rand_bitmap(), rand_bitmap()
        assert a.intersect(b).slice() == sorted(sa & sb)
        assert a.union(b).slice() == sorted(sa | sb)
        assert a.difference(b).slice() == sorted(sa - sb)
        assert a.xor(b).slice() == sorted(sa ^ sb)
        assert a.intersection_count(b) == len(sa & sb)
        (c, sc) = rand_bitmap()
        assert a.union(b, c).slice() == sorted(sa | sb | sc)   # n-way unionInPlace path (roaring.go:737)
