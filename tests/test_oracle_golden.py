"""Pins the CPU oracle against every golden vector / known-answer test the
reference's own test-suite holds for this path."""
import numpy as np
import ctypes as C

_i32p = C.POINTER(C.c_int32)
_u64p = C.POINTER(C.c_uint64)


def _i32(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_i32p)


def test_hamming_kat(oracle):
    # cli/src/sketch.rs:53-58
    a = np.zeros(16, dtype=np.uint64); b = np.zeros(16, dtype=np.uint64)
    a[0] = 0b1111; a[15] = 0b1001; b[15] = 0b1011
    assert oracle.hamming(a, b) == 5


def test_nsw_i32(oracle):
    # src/index/nsw.rs:366-385: points 1..20, defaults m=16 M=32, efc=4; search(&5,4,4) == {3,4,5,6}
    pts, pp = _i32(np.arange(1, 20))
    out = np.zeros(8, dtype=np.int32); od = np.zeros(8, dtype=np.uint64)
    n = oracle.lib.orc_golden_nsw_i32(pp, len(pts), 4, 16, 32, 5, 4, 4, out.ctypes.data_as(_i32p), od.ctypes.data_as(_u64p))
    assert n == 4 and set(out[:4].tolist()) == {3, 4, 5, 6}


def test_hnsw_i32(oracle):
    # src/index/hnsw.rs:309-332: 0..20, m=3, efc=4 (M default 32)
    pts, pp = _i32(np.arange(0, 20))
    for threads in (1, 8, 16):  # the level RNG seed mixes the thread count (hnsw.rs:24-30)
        out = np.zeros(32, dtype=np.int32); od = np.zeros(32, dtype=np.uint64)
        n = oracle.lib.orc_golden_hnsw_i32(pp, 20, 4, 3, 32, threads, 5, 4, 4, out.ctypes.data_as(_i32p), od.ctypes.data_as(_u64p))
        assert n == 4
        assert set(out[:4].tolist()) in ({3, 4, 5, 6}, {4, 5, 6, 7})
        n = oracle.lib.orc_golden_hnsw_i32(pp, 20, 4, 3, 32, threads, 0, 20, 20, out.ctypes.data_as(_i32p), od.ctypes.data_as(_u64p))
        assert n == 20  # whole-index search reaches every node


def test_heuristic(oracle):
    # src/index/nsw.rs:387-412 and src/index/hnsw.rs:334-360
    nums, pn = _i32([1, 5, 6, 7, 16, 18])
    out = np.zeros(8, dtype=np.int32)
    n = oracle.lib.orc_golden_heuristic_i32(pn, 6, 10, 3, out.ctypes.data_as(_i32p))
    assert set(out[:n].tolist()) == {7, 16}


def test_min_k(oracle):
    # src/collections/mod.rs:100-114
    for v, want in (([0, 9, 1, 8, 2, 7, 3, 6, 4, 5, 5, 4, 6, 3, 7, 2, 8, 1, 9, 0], [0, 0, 1, 1, 2]),
                    ([0, 1, 4, 2, 5, 9, 10, 3, 6, 7, 8, 11, 12, 13, 14], [0, 1, 2, 3, 4])):
        a, pa = _i32(v)
        out = np.zeros(5, dtype=np.int32)
        n = oracle.lib.orc_golden_min_k_i32(pa, len(a), 5, out.ctypes.data_as(_i32p))
        assert n == 5 and sorted(out.tolist()) == want
    out = np.zeros(5, dtype=np.int32)
    a, pa = _i32([3, 1, 2])
    assert oracle.lib.orc_golden_min_k_i32(pa, 3, 0, out.ctypes.data_as(_i32p)) == 0  # k == 0 -> []


def test_simplegraph(oracle):
    # src/collections/simplegraph.rs:130-175
    assert oracle.lib.orc_golden_simplegraph() == 0xF


README_EXPECTED = [  # README.md:61-74 / examples/main.rs:53-65
    (49, (2, 4, 9)), (50, (1, 4, 9)), (50, (2, 3, 9)), (50, (2, 5, 9)), (50, (3, 4, 9)),
    (51, (1, 3, 9)), (51, (1, 5, 9)), (51, (3, 3, 9)), (51, (3, 5, 9)), (53, (0, 4, 9)),
]


def _lattice():
    return np.array([(x, y, z) for x in range(10) for y in range(10) for z in range(10)], dtype=np.int32)


def test_readme_example_hnsw(oracle):
    pts = _lattice(); q = np.array([2, 4, 16], dtype=np.int32)
    for threads, parallel in ((1, 0), (8, 0), (16, 0), (16, 1)):
        keys = np.zeros(10, dtype=np.uint64); d = np.zeros(10, dtype=np.uint64)
        n = oracle.lib.orc_golden_hnsw_p3(pts.ctypes.data_as(_i32p), 1000, 24, 8, 32, threads, parallel, q.ctypes.data_as(_i32p), 10, 20,
                                          keys.ctypes.data_as(_u64p), d.ctypes.data_as(_u64p))
        assert n == 10
        got = [(int(d[i]), tuple(int(v) for v in pts[int(keys[i])])) for i in range(10)]
        assert got == README_EXPECTED, (threads, parallel, got)


def test_readme_example_bruteforce(oracle):
    pts = _lattice(); q = np.array([2, 4, 16], dtype=np.int32)
    keys = np.zeros(10, dtype=np.uint64); d = np.zeros(10, dtype=np.uint64)
    n = oracle.lib.orc_golden_bruteforce_p3(pts.ctypes.data_as(_i32p), 1000, q.ctypes.data_as(_i32p), 10,
                                            keys.ctypes.data_as(_u64p), d.ctypes.data_as(_u64p))
    got = [(int(d[i]), tuple(int(v) for v in pts[int(keys[i])])) for i in range(n)]
    assert got == README_EXPECTED
