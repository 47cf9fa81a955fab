"""GPU parity: the CUDA path through the C ABI vs the CPU oracle on the same inputs.
Bit-exact bar (integer work): identical ids, distances, found counts and work counters."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(__file__), "golden")
CASES = [(10, 64), (10, 16), (1, 1), (10, 4), (25, 40), (100, 100)]


@pytest.fixture(scope="module")
def hb():
    import hnsw_itu_b200 as m
    return m


def _queries():
    return np.fromfile(os.path.join(G, "tiny_queries.u64"), dtype=np.uint64).reshape(-1, 16)


def _to_gpu(hb, oidx, keep_host=True):
    return hb.index_from_parts(oidx.algo, oidx.graphs(), oidx.ep, keep_host=keep_host)


def _assert_same(hidx, oidx, q, k, ef, pad=False):
    ok, od, of, oc = oidx.knns(q, k, ef, pad=pad, counters=True)
    gk, gd, gf, gc = hidx.knns_arrays(q, k, ef, pad=pad, counters=True)
    assert np.array_equal(gf, of), (k, ef)
    assert np.array_equal(gk, ok), (k, ef)
    assert np.array_equal(gd, od), (k, ef)
    # X, E, A are determined by the search trajectory and must match; V counts the distance evaluations actually
    # performed: the fast path's recently-seen cache may re-evaluate a visited node (never fewer than |visited|)
    gc = gc.astype(np.uint64)
    assert np.array_equal(gc[:, [0, 1, 3]], oc[:, [0, 1, 3]]), (k, ef)
    assert (gc[:, 2] >= oc[:, 2]).all() and (gc[:, 2] <= oc[:, 1] + 8).all(), (k, ef)


def test_golden_fixtures_from_idx_file(hb):
    """read_index of the committed .idx bytes, then the committed expected ids/distances."""
    exp = np.load(os.path.join(G, "tiny_expected.npz"))
    q = _queries()
    for name in ("hnsw", "nsw"):
        idx = hb.read_index(os.path.join(G, f"tiny_{name}.idx"))
        assert idx.algo == name
        for k, ef in CASES:
            keys, dists, found = idx.knns_arrays(q, k, ef)
            assert np.array_equal(keys, exp[f"{name}_k{k}_ef{ef}_keys"]), (name, k, ef)
            assert np.array_equal(dists, exp[f"{name}_k{k}_ef{ef}_dists"]), (name, k, ef)
            assert np.array_equal(found, exp[f"{name}_k{k}_ef{ef}_found"]), (name, k, ef)
        if name == "hnsw":
            _, _, _, ctr = idx.knns_arrays(q, 10, 64, counters=True)
            ctr = ctr.astype(np.uint64)
            assert np.array_equal(ctr[:, [0, 1, 3]], exp["hnsw_k10_ef64_ctr"][:, [0, 1, 3]])
            assert (ctr[:, 2] >= exp["hnsw_k10_ef64_ctr"][:, 2]).all()
    bf = hb.read_index(os.path.join(G, "tiny_bruteforce.idx"))
    assert bf.algo == "bruteforce" and bf.size() == 100


def test_idx_write_roundtrip(hb, oracle, tmp_path):
    import filecmp
    for name in ("hnsw", "nsw", "bruteforce"):
        src = os.path.join(G, f"tiny_{name}.idx")
        dst = str(tmp_path / f"{name}.idx")
        hb.read_index(src).write_index(dst)
        assert filecmp.cmp(src, dst, shallow=False)


@pytest.mark.parametrize("distinct_planes", [1024, 64, 16])
def test_search_parity_vs_oracle(hb, oracle, distinct_planes):
    from oracle import oracle as orc
    p = orc.SynthParams(latent_dim=16, n_clusters=64, centre_scale=1, noise_scale=1, distinct_planes=distinct_planes)
    data = oracle.synth(6000, 31, p)
    q = oracle.synth(300, 32, p)
    oidx = oracle.build(data, "hnsw", efc=32, m=8, M=24, parallel=False, rayon_threads=4)
    hidx = _to_gpu(hb, oidx)
    assert hidx.size() == 6000
    for ef in (0, 1, 2, 10, 16, 33, 64, 96, 256):
        for k in (1, 10, 64, 100):
            _assert_same(hidx, oidx, q, k, ef)
    _assert_same(hidx, oidx, q, 10, 4, pad=True)


def test_search_parity_reference_parameters(hb, oracle):
    """m=24 M=256 c=96 (the reference's CLI defaults) incl. hub rows longer than one 128-B line."""
    data = oracle.synth(20000, 41)
    q = oracle.synth(500, 42)
    oidx = oracle.build(data, "hnsw", efc=96, m=24, M=256, parallel=False)
    hidx = _to_gpu(hb, oidx, keep_host=False)
    for ef in (16, 64, 96, 128):
        _assert_same(hidx, oidx, q, 10, ef)
    # Index::search == knns of one
    res = hidx.search(q[0], 10, 64)
    exp = oidx.search(q[0], 10, 64)
    assert [(d.distance, d.key) for d in res] == exp
    assert res == sorted(res)


def test_insert_at_last_beam_slot_regression(hb, oracle):
    """Inputs that once exposed a missing warp barrier in beam_insert (insert at position ef-1 of a full beam):
    some lanes then saw the new element as the evicted one and the warp's state diverged."""
    data = oracle.synth(20000, 1003)
    q = oracle.synth(2000, 1004)
    oidx = oracle.build(data, "hnsw", efc=96, m=24, M=256, parallel=False)
    hidx = _to_gpu(hb, oidx, keep_host=False)
    for ef in (16, 64):
        _assert_same(hidx, oidx, q, 10, ef)


def test_nsw_and_edge_cases(hb, oracle):
    from oracle import oracle as orc
    p = orc.SynthParams(latent_dim=8, n_clusters=16, distinct_planes=32)
    data = oracle.synth(1500, 21, p)
    q = oracle.synth(64, 22, p)
    onsw = oracle.build(data, "nsw", efc=16, m=4, M=8, parallel=False)
    hnsw = _to_gpu(hb, onsw)
    for ef in (1, 5, 40):
        _assert_same(hnsw, onsw, q, 10, ef)
    # single node (self loop), fewer nodes than k, padding like the CLI
    one = oracle.build(data[:1], "hnsw", efc=8, m=4, M=8, parallel=False)
    h1 = _to_gpu(hb, one)
    _assert_same(h1, one, q[:3], 10, 64)
    _assert_same(h1, one, q[:3], 10, 64, pad=True)
    # empty index: HNSW::search returns [] (hnsw.rs:282); padding then fails like the reference's panic
    empty = hb.index_from_parts("hnsw", [{"sketches": np.zeros((0, 16), dtype=np.uint64), "off": np.zeros(1, dtype=np.uint64),
                                          "ids": np.zeros(0, dtype=np.uint64)}], None)
    keys, dists, found = empty.knns_arrays(q[:2], 5, 10)
    assert (found == 0).all() and (keys == 0).all()
    with pytest.raises(hb.HbError) as e:
        empty.knns_arrays(q[:2], 5, 10, pad=True)
    assert e.value.code == 7
    # zero queries
    k0, d0, f0 = hnsw.knns_arrays(np.zeros((0, 16), dtype=np.uint64), 5, 10)
    assert k0.shape == (0, 5)


def test_visited_table_overflow_retry(hb, oracle):
    """ef far above what the first visited-table size allows for: forces the exact-overflow retry path."""
    from oracle import oracle as orc
    p = orc.SynthParams(latent_dim=8, n_clusters=4, distinct_planes=1024)
    data = oracle.synth(30000, 51, p)
    q = oracle.synth(40, 52, p)
    oidx = oracle.build(data, "hnsw", efc=32, m=16, M=256, parallel=False)
    hidx = _to_gpu(hb, oidx, keep_host=False)
    _assert_same(hidx, oidx, q, 10, 16)   # table starts at 4096 slots; hub-heavy graph overflows it for some queries
    _assert_same(hidx, oidx, q, 100, 1000)


def test_ground_truth_parity(hb, oracle):
    from oracle import oracle as orc
    for dp, n, nq, k in ((1024, 5000, 70, 10), (32, 3000, 33, 100), (1024, 7, 5, 10), (1024, 1, 3, 1)):
        p = orc.SynthParams(latent_dim=8, n_clusters=16, distinct_planes=dp)
        data = oracle.synth(n, 61, p)
        q = oracle.synth(nq, 62, p)
        ok, od = oracle.ground_truth(data, q, k)
        m = min(n, k)
        for algo in ("auto", "scan", "mma", "mxf4"):
            gk, gd = hb.ground_truth(data, q, k, algo=algo)
            assert np.array_equal(gk[:, :m], ok[:, :m]) and np.array_equal(gd[:, :m], od[:, :m]), (dp, n, nq, k, algo)
            if n < k:
                assert (gk[:, m:] == np.uint64(2**64 - 1)).all()
    exp = np.load(os.path.join(G, "tiny_expected.npz"))
    data = oracle.load_idx(os.path.join(G, "tiny_hnsw.idx")).graph(0)["sketches"]
    gk, gd = hb.ground_truth(data, _queries(), 10)
    assert np.array_equal(gk, exp["gt_k10_keys"]) and np.array_equal(gd, exp["gt_k10_dists"])
    # Indexes::Bruteforce through Index::knns (sorted order; ef ignored)
    bf = hb.read_index(os.path.join(G, "tiny_bruteforce.idx"))
    bk, bd, bfnd = bf.knns_arrays(_queries(), 10, 3)
    ok, od = oracle.ground_truth(data[:100], _queries(), 10)
    assert np.array_equal(bk, ok) and np.array_equal(bd, od) and (bfnd == 10).all()


@pytest.mark.parametrize("dp,n,nq,k", [
    (1024, 20000, 300, 10),   # several query tiles (last one ragged), many data tiles
    (16, 9001, 129, 100),     # distances in multiples of 64: ties everywhere, k = 100, ragged last data tile
    (1024, 255, 1, 10),       # less than one data tile, one query
    (64, 70000, 128, 1),      # k = 1; more items than SMs' worth of chunks
    (1024, 513, 257, 1024),   # k > rows in most chunks, k at the supported maximum
])
@pytest.mark.parametrize("algo", ["mma", "mxf4"])
def test_ground_truth_tcgen05_vs_oracle(hb, oracle, dp, n, nq, k, algo):
    """The tcgen05 GEMM formulations (int8: gt_tcgen05.cu, E2M1: gt_mxf4.cu) against Bruteforce::search + sort restated by the oracle,
    and against the XOR+POPC scan kernel: ids and distances identical, including the (distance, key) tie order."""
    from oracle import oracle as orc
    p = orc.SynthParams(latent_dim=8, n_clusters=16, distinct_planes=dp)
    data = oracle.synth(n, 161, p)
    q = oracle.synth(nq, 162, p)
    q[0] = data[n // 2]            # a query that is a member: distance 0
    if nq > 1:
        q[-1] = ~data[n - 1]       # and the complement of the last row: distance 1024 to it
    mk, md = hb.ground_truth(data, q, k, algo=algo)
    sk, sd = hb.ground_truth(data, q, k, algo="scan")
    assert np.array_equal(mk, sk) and np.array_equal(md, sd)
    ok, od = oracle.ground_truth(data, q, k)
    m = min(n, k)
    assert np.array_equal(mk[:, :m], ok[:, :m]) and np.array_equal(md[:, :m], od[:, :m])
    assert md[0, 0] == 0 and mk[0, 0] <= n // 2


@pytest.mark.parametrize("plan", [1, 2])
@pytest.mark.parametrize("dp,n,nq,k", [
    (1024, 20000, 300, 10),   # three query tiles: a full pair and a pair whose second CTA has no queries
    (16, 9001, 257, 100),     # ties everywhere, k = 100, ragged last data tile, odd number of query tiles
    (1024, 119, 129, 10),     # less than one 120-row tile
    (1024, 513, 257, 1024),   # k at the supported maximum
])
def test_ground_truth_mxf4_kernel_plans(hb, oracle, monkeypatch, plan, dp, n, nq, k):
    """The E2M1 kernel's alternative plans (HB_GT_PLAN: 1 = CTA pairs issuing tcgen05.mma.cta_group::2 with each CTA
    expanding half of every data tile, 2 = four 120-row accumulators) against the default plan and the oracle."""
    from oracle import oracle as orc
    p = orc.SynthParams(latent_dim=8, n_clusters=16, distinct_planes=dp)
    data = oracle.synth(n, 171, p)
    q = oracle.synth(nq, 172, p)
    q[0] = data[n // 2]
    q[-1] = ~data[n - 1]
    monkeypatch.delenv("HB_GT_PLAN", raising=False)
    dk, dd = hb.ground_truth(data, q, k, algo="mxf4")
    monkeypatch.setenv("HB_GT_PLAN", str(plan))
    mk, md = hb.ground_truth(data, q, k, algo="mxf4")
    monkeypatch.delenv("HB_GT_PLAN", raising=False)
    assert np.array_equal(mk, dk) and np.array_equal(md, dd)
    ok, od = oracle.ground_truth(data, q, k)
    m = min(n, k)
    assert np.array_equal(mk[:, :m], ok[:, :m]) and np.array_equal(md[:, :m], od[:, :m])


@pytest.mark.parametrize("algo", ["mma", "mxf4"])
def test_ground_truth_tcgen05_all_bit_positions(hb, algo):
    """Every one of the 1024 bit positions contributes exactly 1 to the distance (the operand scaling/permutation
    of the int8 / E2M1 expansion is position dependent): rows = one-hot sketches against the zero query and its complement."""
    data = np.zeros((1024 + 300, 16), dtype=np.uint64)
    for b in range(1024):
        data[b, b // 64] = np.uint64(1) << np.uint64(b % 64)
    for r in range(300):   # r low bits set
        bits = r + 2
        for w in range(16):
            nb = min(64, max(0, bits - 64 * w))
            data[1024 + r, w] = np.uint64((1 << nb) - 1) if nb < 64 else np.uint64(2**64 - 1)
    q = np.zeros((40, 16), dtype=np.uint64)
    q[1] = ~q[1]
    for i in range(2, 40):
        q[i] = data[(i * 37) % 1024] | data[(i * 91 + 5) % 1024]
    mk, md = hb.ground_truth(data, q, 50, algo=algo)
    sk, sd = hb.ground_truth(data, q, 50, algo="scan")
    assert np.array_equal(mk, sk) and np.array_equal(md, sd)
    assert (md[0, :50] == 1).all() and np.array_equal(mk[0, :50], np.arange(50, dtype=np.uint64))
    x = data[:, None, :] ^ q[None, :3, :]
    ref = np.unpackbits(x.view(np.uint8), axis=-1).sum(-1).astype(np.uint64)   # (n, 3)
    for j in range(3):
        order = np.lexsort((np.arange(len(data)), ref[:, j]))[:50]
        assert np.array_equal(mk[j], order.astype(np.uint64)) and np.array_equal(md[j], ref[order, j])


def test_sharded_ground_truth_merge(hb, oracle):
    """config 4's shape on one GPU: per-shard top-k with key offsets, then the exact (distance,key) merge."""
    import torch
    from hnsw_itu_b200 import _lib
    data = oracle.synth(4000, 71)
    q = oracle.synth(50, 72)
    k, shards = 10, 4
    dq = torch.from_numpy(q.view(np.int64)).cuda()
    lists = torch.empty((shards, len(q), k), dtype=torch.int64, device="cuda")
    for s in range(shards):
        part = torch.from_numpy(data[s * 1000:(s + 1) * 1000].view(np.int64)).cuda()
        _lib.check(_lib.lib().hb_ground_truth_device(part.data_ptr(), 1000, s * 1000, dq.data_ptr(), len(q), k, lists[s].data_ptr(), None))
    merged = torch.empty((len(q), k), dtype=torch.int64, device="cuda")
    _lib.check(_lib.lib().hb_topk_merge_device(lists.data_ptr(), shards, len(q), k, merged.data_ptr(), None))
    keys = torch.empty_like(merged); dists = torch.empty_like(merged)
    _lib.check(_lib.lib().hb_unpack_device(merged.data_ptr(), merged.numel(), keys.data_ptr(), dists.data_ptr(), None))
    torch.cuda.synchronize()
    ok, od = oracle.ground_truth(data, q, k)
    assert np.array_equal(keys.cpu().numpy().view(np.uint64), ok)
    assert np.array_equal(dists.cpu().numpy().view(np.uint64), od)


def test_synth_kernel_matches_oracle(hb, oracle):
    import torch
    from oracle import oracle as orc
    for p in (orc.SynthParams(), orc.SynthParams(latent_dim=12, n_clusters=24, distinct_planes=128)):
        n, start = 1000, 12345
        out = torch.empty((n, 16), dtype=torch.int64, device="cuda")
        hb.synth_sketches(out.data_ptr(), n, 1003, start=start, latent_dim=p.latent_dim, n_clusters=p.n_clusters,
                          centre_scale=p.centre_scale, noise_scale=p.noise_scale, distinct_planes=p.distinct_planes)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy().view(np.uint64), oracle.synth(n, 1003, p, start=start))


def test_size_independent_properties(hb, oracle):
    """Properties that hold at any size: ascending (distance,key), no duplicate keys, distances recomputable,
    self-query finds itself at distance 0, ground truth of a permuted shard union is unchanged."""
    data = oracle.synth(50000, 81)
    oidx = oracle.build(data[:20000], "hnsw", efc=48, m=12, M=64, parallel=False)
    hidx = _to_gpu(hb, oidx, keep_host=False)
    q = data[:256]
    keys, dists, found = hidx.knns_arrays(q, 10, 64)
    assert (found == 10).all()
    assert (keys[:, 0] == np.arange(256)).all() and (dists[:, 0] == 0).all()
    packed = (dists.astype(np.uint64) << np.uint64(32)) | keys
    assert (np.diff(packed.astype(np.int64), axis=1) > 0).all()
    for i in range(0, 256, 37):
        for j in range(10):
            assert oracle.hamming(q[i], data[int(keys[i, j])]) == int(dists[i, j])
    gk, gd = hb.ground_truth(data, q[:32], 10)
    assert (gk[:, 0] == np.arange(32)).all() and (gd[:, 0] == 0).all()
    assert (np.diff(((gd << np.uint64(32)) | gk).astype(np.int64), axis=1) > 0).all()


def test_sharded_hnsw_pack_and_merge(hb, oracle):
    """config 5's shape on one GPU: one index per row range (the reference's `index -b start -l len`), same ef on every
    shard, local keys shifted by the shard offset (hb_pack_device), exact merge == merge of the oracle's per-shard lists.
    k > found on small shards exercises the "no element" padding."""
    import torch
    from hnsw_itu_b200 import sharded
    data = oracle.synth(9000, 91)
    q = oracle.synth(120, 92)
    k, ef = 10, 24
    bounds = sharded.shard_bounds(len(data), 3)
    bounds[-1] = (bounds[-1][0], 6)          # a shard with fewer than k rows
    dq = torch.from_numpy(q.view(np.int64)).cuda()
    lists = torch.empty((len(bounds), len(q), k), dtype=torch.int64, device="cuda")
    expect = []
    for s, (start, ln) in enumerate(bounds):
        oidx = oracle.build(data[start:start + ln], "hnsw", efc=32, m=8, M=24, parallel=False)
        hidx = _to_gpu(hb, oidx, keep_host=False)
        lists[s] = sharded.knns_sharded(hidx, start, dq, k, ef)   # world 1: pack + merge of a single list
        ok, od, of = oidx.knns(q, k, ef)
        pk = (od.astype(np.uint64) << np.uint64(32)) | (ok + np.uint64(start))
        pk[np.arange(k)[None, :] >= of[:, None]] = np.uint64(2**64 - 1)
        expect.append(pk)
        assert np.array_equal(lists[s].cpu().numpy().view(np.uint64), pk)
    merged = sharded.merge_lists_device(lists)
    keys, dists = sharded.unpack_device(merged)
    torch.cuda.synchronize()
    want = np.sort(np.concatenate(expect, axis=1), axis=1)[:, :k]
    assert np.array_equal(merged.cpu().numpy().view(np.uint64), want)
    assert np.array_equal(keys.cpu().numpy().view(np.uint64), want & np.uint64(0xFFFFFFFF))
    # sharded ground truth through the same host function (world 1)
    dd = torch.from_numpy(data.view(np.int64)).cuda()
    gt = sharded.ground_truth_sharded(dd, 0, dq, k)
    gk, gd = oracle.ground_truth(data, q, k)
    assert np.array_equal(gt.cpu().numpy().view(np.uint64), (gd << np.uint64(32)) | gk)


@pytest.mark.parametrize("k", [10, 100])
def test_ground_truth_large_properties(hb, k):
    """Size-independent checks at a size the oracle would not finish in seconds (2 M x 1 k, generated on the device):
    both tcgen05 kernels == scan kernel bit for bit; rows ascend in (distance, key) with unique keys; every returned distance
    is recomputed from the sketches; queries that are data rows find themselves at distance 0."""
    import torch
    from hnsw_itu_b200 import _lib
    n, nq = 2_000_000, 1024
    d = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    q = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(d.data_ptr(), n, 1003)
    hb.synth_sketches(q.data_ptr(), nq, 1004)
    members = torch.arange(0, 64, device="cuda") * 31_013 + 7          # 64 queries are copies of data rows
    q[:64] = d[members]
    out = {}
    for algo in (3, 2, 1):                                              # E2M1 GEMM, int8 GEMM, XOR+POPC scan
        p = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        _lib.check(_lib.lib().hb_ground_truth_device_algo(d.data_ptr(), n, 0, q.data_ptr(), nq, k, algo, p.data_ptr(), None))
        torch.cuda.synchronize()
        out[algo] = p.cpu().numpy().view(np.uint64)
    assert np.array_equal(out[1], out[2]) and np.array_equal(out[1], out[3])
    packed = out[3]
    assert (packed[:, 1:] > packed[:, :-1]).all()                       # strictly ascending (distance << 32 | key): unique keys
    keys, dists = (packed & np.uint64(0xFFFFFFFF)).astype(np.int64), (packed >> np.uint64(32)).astype(np.int64)
    assert keys.max() < n
    hd, hq = d.cpu().numpy().view(np.uint64), q.cpu().numpy().view(np.uint64)
    rows = np.arange(0, nq, 7)
    x = hd[keys[rows]] ^ hq[rows][:, None, :]                           # (r, k, 16)
    ref = np.unpackbits(x.view(np.uint8), axis=-1).sum(-1)
    assert np.array_equal(ref, dists[rows])
    assert (dists[:64, 0] == 0).all() and (keys[:64, 0] <= members.cpu().numpy()).all()


@pytest.mark.parametrize("k", [10, 100, 1000])
def test_ground_truth_all_rows_tied(hb, k):
    """Every row at the same distance from every query (identical rows, plus a block of rows one bit closer in the
    middle): the top-k is decided by the keys alone, the per-row thresholds never separate anything, and the append
    buffers fill with ties (the compaction's key-bisection fallback).  Expected lists are known in closed form."""
    n, nq = 6000, 40
    base = np.full(16, 0x5555555555555555, dtype=np.uint64)
    data = np.tile(base, (n, 1))
    closer = np.arange(3000, 3000 + 7)                      # seven rows that differ from the rest in bit 0 of word 0
    data[closer, 0] ^= np.uint64(1)
    q = np.tile(base, (nq, 1))
    q[:, 0] ^= np.uint64(1)                                 # queries equal the `closer` rows: distance 0 to them, 1 to the rest
    q[1::2, 15] ^= np.uint64(0xFF00)                        # odd queries 8 bits further from everything
    exp_keys = np.concatenate([closer, np.setdiff1d(np.arange(n), closer)])[:k].astype(np.uint64)
    for algo in ("mxf4", "mma", "scan"):
        gk, gd = hb.ground_truth(data, q, k, algo=algo)
        assert np.array_equal(gk, np.tile(exp_keys, (nq, 1))), algo
        d0 = np.where(np.arange(k) < 7, 0, 1).astype(np.uint64)
        assert np.array_equal(gd[0::2], np.tile(d0, (nq // 2, 1))) and np.array_equal(gd[1::2], np.tile(d0 + np.uint64(8), (nq // 2, 1))), algo


def test_ground_truth_host_buffers_sliced_upload(hb):
    """hb_ground_truth on host buffers above the slice size (2^21 rows): the rows are uploaded in slices while the
    previous slice is searched, per-slice lists are merged exactly -- same lists as one search over the resident rows."""
    import torch
    from hnsw_itu_b200 import _lib
    n, nq, k = 4_500_000, 256, 10                                       # three slices, the last one ragged
    d = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    q = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(d.data_ptr(), n, 2003)
    hb.synth_sketches(q.data_ptr(), nq, 2004)
    q[:8] = d[torch.tensor([0, 2_097_151, 2_097_152, 4_194_303, 4_194_304, 4_499_999, 17, 3_000_000], device="cuda")]  # slice borders
    p = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    _lib.check(_lib.lib().hb_ground_truth_device(d.data_ptr(), n, 0, q.data_ptr(), nq, k, p.data_ptr(), None))
    torch.cuda.synchronize()
    ref = p.cpu().numpy().view(np.uint64)
    hd, hq = d.cpu().numpy().view(np.uint64), q.cpu().numpy().view(np.uint64)
    del d, p
    keys, dists = hb.ground_truth(hd, hq, k)
    assert np.array_equal((dists << np.uint64(32)) | keys, ref)
    assert (dists[:8, 0] == 0).all()


def test_knns_pinned_buffers_used_in_place(hb, oracle):
    """hb_knns with page-locked caller buffers (what a serving loop passes): the kernel reads the queries and writes
    the results in host memory directly; same ids, distances and found counts as the staged path and the oracle."""
    import ctypes as C
    import torch
    from hnsw_itu_b200 import _lib
    data = oracle.synth(20000, 81)
    q = oracle.synth(300, 82)
    oidx = oracle.build(data, "hnsw", efc=48, m=12, M=64, parallel=False)
    hidx = _to_gpu(hb, oidx, keep_host=False)
    k, ef = 10, 64
    hq = torch.from_numpy(q.view(np.int64)).pin_memory()
    hk = torch.full((len(q), k), -1, dtype=torch.int64).pin_memory()
    hd = torch.full((len(q), k), -1, dtype=torch.int64).pin_memory()
    hf = torch.full((len(q),), -1, dtype=torch.int32).pin_memory()
    u64p, u32p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
    for _ in range(2):
        _lib.check(_lib.lib().hb_knns(hidx._h, C.cast(hq.data_ptr(), u64p), len(q), k, ef, 0, C.cast(hk.data_ptr(), u64p),
                                      C.cast(hd.data_ptr(), u64p), C.cast(hf.data_ptr(), u32p), None))
    ok, od, of = oidx.knns(q, k, ef)
    assert np.array_equal(hk.numpy().view(np.uint64), ok) and np.array_equal(hd.numpy().view(np.uint64), od)
    assert np.array_equal(hf.numpy().view(np.uint32), of)
    gk, gd, gf = hidx.knns_arrays(q, k, ef)                              # pageable numpy buffers: staged copies
    assert np.array_equal(gk, ok) and np.array_equal(gd, od) and np.array_equal(gf, of)


def test_compact_layout_returns_the_same_rows(hb, oracle):
    """hb_index_compact: line-granular adjacency rows + offsets instead of fixed-stride rows -- a layout change only."""
    data = oracle.synth(20000, 91)
    q = oracle.synth(300, 92)
    oidx = oracle.build(data, "hnsw", efc=48, m=12, M=96, parallel=False)   # hub rows longer than one two-line segment
    hidx = _to_gpu(hb, oidx)
    before = hidx.info().device_bytes
    assert hidx.info().row_stride >= 64
    hidx.compact()
    inf = hidx.info()
    assert inf.row_stride == 0 and inf.device_bytes < before
    for k, ef in ((10, 64), (10, 16), (100, 256), (10, 300)):      # both kernels (ef <= 256 and the large-capacity one)
        _assert_same(hidx, oidx, q, k, ef)
    hidx.compact()                                                    # idempotent
    _assert_same(hidx, oidx, q, 10, 64)
    # a built index compacts too
    b = hb.build_hnsw(data, efc=48, m=12, M=48, serial_prefix=2000)
    k0, d0, f0 = b.knns_arrays(q, 10, 64)
    b.compact()
    k1, d1, f1 = b.knns_arrays(q, 10, 64)
    assert np.array_equal(k0, k1) and np.array_equal(d0, d1) and np.array_equal(f0, f1)


def test_parity_at_one_million_points(hb, oracle):
    """Full-size behaviour inside -m gpu: a 1 M-point index (device builder, reference parameters), every query of a
    2 000-query batch identical to the oracle's on the same graph (ids, distances, X/E/A), plus the large-capacity path
    (ef = 1000: exact visited tables sized per call on the caller's stream) on a 10 k batch through hb_knns."""
    import torch
    n, nq = 1_000_000, 10_000
    d = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(d.data_ptr(), n, 1003)
    dq = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(dq.data_ptr(), nq, 1004)
    torch.cuda.synchronize()
    data, q = d.cpu().numpy().view(np.uint64), dq.cpu().numpy().view(np.uint64)
    del d, dq
    hidx = hb.build_hnsw(data, efc=96, m=24, M=256, rayon_threads=16, keep_host=True)
    inf = hidx.info()
    assert inf.size == n and inf.build_dropped_appends == 0
    oidx = oracle.from_parts("hnsw", [hidx.graph(g) for g in range(inf.n_graphs)], inf.entry_point)
    _assert_same(hidx, oidx, q[:2000], 10, 64)
    _assert_same(hidx, oidx, q[:500], 10, 256)
    # ef = 1000 over the whole batch on pageable host buffers: first use of the visited tables (memset on the call's stream)
    gk, gd, gf = hidx.knns_arrays(q, 10, 1000)
    ok, od, of = oidx.knns(q[:300], 10, 1000)
    assert np.array_equal(gk[:300], ok) and np.array_equal(gd[:300], od) and np.array_equal(gf[:300], of)
    assert (gf == 10).all() and (np.diff(gd.astype(np.int64), axis=1) >= 0).all()
    # and again after compaction of the adjacency
    hidx.compact()
    _assert_same(hidx, oidx, q[:500], 10, 64)


def test_exact_visited_flag_reproduces_the_references_visited_count(hb, oracle):
    """HB_EXACT_VISITED: the large-capacity kernel at any ef; all four counters (X, E, V, A) equal the oracle's."""
    data, q = oracle.synth(8000, 93), oracle.synth(200, 94)
    oidx = oracle.build(data, "hnsw", efc=48, m=12, M=48, parallel=False)
    hidx = _to_gpu(hb, oidx)
    for k, ef in ((10, 64), (10, 16)):
        ok, od, of, oc = oidx.knns(q, k, ef, counters=True)
        gk, gd, gf, gc = hidx.knns_arrays(q, k, ef, counters=True, exact_visited=True)
        assert np.array_equal(gk, ok) and np.array_equal(gd, od) and np.array_equal(gf, of)
        assert np.array_equal(gc.astype(np.uint64), oc), (k, ef)
