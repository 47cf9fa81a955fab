"""hb_build_hnsw (index construction on the device).  The built graph cannot equal the reference's
(HashSet order, rayon thread count in the seed), so the checks are structural (set semantics, symmetry, degree
bound), statistical (recall next to the CPU restatement of the builder) and -- the parity bar of the search path --
bit-exact agreement of GPU search and oracle search on the graph the device built."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hb():
    import hnsw_itu_b200 as m
    return m


def _check_structure(idx, M, max_new_per_step):
    info = idx.info()
    assert info.build_dropped_appends == 0
    graphs = [idx.graph(g) for g in range(info.n_graphs)]
    for gi, g in enumerate(graphs):
        n = len(g["sketches"])
        off, ids = g["off"].astype(np.int64), g["ids"].astype(np.int64)
        assert off[-1] == len(ids) and (ids >= 0).all() and (ids < n).all()
        deg = np.diff(off)
        assert deg.max(initial=0) <= M + max_new_per_step, (gi, deg.max())
        src = np.repeat(np.arange(n), deg)
        fwd = src * n + ids
        assert len(np.unique(fwd)) == len(fwd), "duplicate edge"
        assert np.array_equal(np.sort(fwd), np.sort(ids * n + src)), "adjacency is not symmetric"
        if gi > 0:
            below = graphs[gi - 1]["sketches"]
            assert np.array_equal(below[g["down"].astype(np.int64)], g["sketches"]), "down-link does not reach the same sketch"
    return graphs


@pytest.mark.parametrize("cfg", [dict(n=3000, prefix=0, threads=16, efc=32, m=8, M=32),      # all through `add`
                                 dict(n=9000, prefix=1000, threads=2, efc=48, m=12, M=64),   # chunks of 64
                                 dict(n=6000, prefix=500, threads=4, efc=32, m=8, M=12)])    # small M: pruning + unlinking
def test_build_then_search_parity(hb, oracle, cfg):
    from oracle import oracle as orc
    data = oracle.synth(cfg["n"], 91)
    q = oracle.synth(300, 92)
    idx = hb.build_hnsw(data, efc=cfg["efc"], m=cfg["m"], M=cfg["M"], rayon_threads=cfg["threads"], serial_prefix=cfg["prefix"])
    assert idx.size() == cfg["n"] and idx.algo == "hnsw"
    step = 1 if cfg["prefix"] == 0 else 32 * cfg["threads"]
    graphs = _check_structure(idx, cfg["M"], step)
    if cfg["M"] == 12:
        assert idx.info().build_prunes > 0 and idx.info().build_unlinked > 0
    oidx = oracle.from_parts("hnsw", graphs, idx.info().entry_point)
    for ef in (1, 10, 64):
        gk, gd, gf, gc = idx.knns_arrays(q, 10, ef, counters=True)
        ok, od, of, oc = oidx.knns(q, 10, ef, counters=True)
        assert np.array_equal(gk, ok) and np.array_equal(gd, od) and np.array_equal(gf, of)
        assert np.array_equal(gc.astype(np.uint64)[:, [0, 1, 3]], oc[:, [0, 1, 3]])
    # quality next to the CPU restatement of the reference's builder, same parameters
    gt, _ = oracle.ground_truth(data, q, 10)
    gk, _, _ = idx.knns_arrays(q, 10, 64)
    cpu = oracle.build(data, "hnsw", efc=cfg["efc"], m=cfg["m"], M=cfg["M"], parallel=cfg["prefix"] != 0, rayon_threads=cfg["threads"])
    ck, _, _ = cpu.knns(q, 10, 64)
    r_gpu, r_cpu = orc.recall_at_k(gk, gt, 10), orc.recall_at_k(ck, gt, 10)
    assert r_gpu >= r_cpu - 0.03, (r_gpu, r_cpu)


def test_built_index_roundtrips_through_idx_file(hb, oracle, tmp_path):
    data = oracle.synth(4000, 93)
    q = oracle.synth(100, 94)
    idx = hb.build_hnsw(data, efc=32, m=8, M=32, rayon_threads=4, serial_prefix=800)
    path = str(tmp_path / "built.idx")
    idx.write_index(path)
    oidx = oracle.load_idx(path)     # the reference's bincode layout, parsed by the oracle's reader
    back = hb.read_index(path)
    a = idx.knns_arrays(q, 10, 48)
    b = back.knns_arrays(q, 10, 48)
    c = oidx.knns(q, 10, 48)
    for x, y, z in zip(a, b, c):
        assert np.array_equal(x, y) and np.array_equal(x, z)


def test_build_edge_cases(hb, oracle):
    data = oracle.synth(40, 95)
    one = hb.build_hnsw(data[:1], efc=8, m=4, M=8)
    g = one.graph(0)
    assert list(g["ids"]) == [0]            # the first point links to itself (hnsw.rs:185-190, 235-243)
    k, d, f = one.knns_arrays(data[:2], 5, 8)
    assert f.tolist() == [1, 1] and k[0, 0] == 0 and d[0, 0] == 0
    empty = hb.build_hnsw(np.zeros((0, 16), dtype=np.uint64))
    assert empty.size() == 0
    small = hb.build_hnsw(data, efc=8, m=4, M=8)
    k, d, f = small.knns_arrays(data, 1, 16)
    assert (k[:, 0] == np.arange(40)).all() and (d[:, 0] == 0).all()
