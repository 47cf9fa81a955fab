"""GPU: the multi-device entry points of the C ABI (csrc/multi.cu: hb_multi_*, hb_comm_*), checked against the
single-device path and the oracle.  The one-device cases run on any box (a communicator of one rank: no NCCL call);
the two-device cases need a box with >= 2 GPUs (`gpurun --gpus 2`) and exercise ncclBroadcast / ncclAllGather inside
the library from one process."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hb():
    import hnsw_itu_b200 as m
    return m


def _n_devices():
    from hnsw_itu_b200 import _lib
    n = C.c_int(0)
    _lib.check(_lib.lib().hb_device_count(C.byref(n)))
    return n.value


def _shards(oracle, data, world, tmp_path):
    """The reference's `index -b start -l len` per shard, built by the oracle and written as .idx files."""
    n = len(data)
    paths, offs, oidxs = [], [], []
    for r in range(world):
        s, e = r * n // world, (r + 1) * n // world
        o = oracle.build(data[s:e], "hnsw", efc=32, m=8, M=24, parallel=False)
        p = str(tmp_path / f"shard{r}.idx")
        o.save_idx(p)
        paths.append(p)
        offs.append(s)
        oidxs.append(o)
    return paths, offs, oidxs


def _merged_oracle(oidxs, offs, q, k, ef):
    """Per-shard oracle lists shifted to global keys and merged by (distance, key): what the sharded search must return."""
    lists = []
    for o, off in zip(oidxs, offs):
        ok, od, of = o.knns(q, k, ef)
        packed = (od << np.uint64(32)) | (ok + np.uint64(off))
        packed[np.arange(k)[None, :] >= of[:, None]] = np.uint64(0xFFFFFFFFFFFFFFFF)
        lists.append(packed)
    m = np.sort(np.concatenate(lists, axis=1), axis=1)[:, :k]
    return m


@pytest.mark.parametrize("world", [1, 2])
def test_replicated_knns_equals_single_device(hb, oracle, world):
    if _n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    data = oracle.synth(5000, 61)
    q = oracle.synth(257, 62)   # odd batch: uneven slices
    oidx = oracle.build(data, "hnsw", efc=32, m=8, M=24, parallel=False)
    one = hb.index_from_parts("hnsw", oidx.graphs(), oidx.ep, device=0)
    multi = hb.MultiIndex(list(range(world))).replicate(one)
    assert multi.size() == 5000
    for k, ef, pad in ((10, 64, False), (10, 4, True), (1, 1, False)):
        ek, ed, ef_ = oidx.knns(q, k, ef, pad=pad)
        gk, gd, gf = multi.knns_arrays(q, k, ef, pad=pad)
        assert np.array_equal(gk, ek) and np.array_equal(gd, ed) and np.array_equal(gf, ef_), (k, ef, pad)


@pytest.mark.parametrize("world", [1, 2])
def test_load_idx_replicates_the_file(hb, oracle, world, tmp_path):
    if _n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    data = oracle.synth(3000, 63)
    q = oracle.synth(100, 64)
    oidx = oracle.build(data, "hnsw", efc=32, m=8, M=24, parallel=False)
    path = str(tmp_path / "x.idx")
    oidx.save_idx(path)
    multi = hb.MultiIndex(list(range(world))).load_idx(path)
    ek, ed, ef_ = oidx.knns(q, 10, 64)
    gk, gd, gf = multi.knns_arrays(q, 10, 64)
    assert np.array_equal(gk, ek) and np.array_equal(gd, ed) and np.array_equal(gf, ef_)


@pytest.mark.parametrize("world", [1, 2])
def test_sharded_hnsw_equals_merged_per_shard_oracle(hb, oracle, world, tmp_path):
    if _n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    data = oracle.synth(6001, 65)
    q = oracle.synth(200, 66)
    paths, offs, oidxs = _shards(oracle, data, world, tmp_path)
    multi = hb.MultiIndex(list(range(world))).load_shards(paths, offs)
    assert multi.size() == 6001
    for k, ef in ((10, 64), (10, 4)):
        exp = _merged_oracle(oidxs, offs, q, k, ef)
        gk, gd, gf = multi.knns_arrays(q, k, ef)
        present = np.arange(k)[None, :] < gf[:, None]
        got = np.where(present, (gd << np.uint64(32)) | gk, np.uint64(0xFFFFFFFFFFFFFFFF))
        assert np.array_equal(got, exp), (k, ef)
        assert np.array_equal(gf, (exp != np.uint64(0xFFFFFFFFFFFFFFFF)).sum(axis=1).astype(np.uint32))


@pytest.mark.parametrize("world", [1, 2])
def test_sharded_ground_truth_equals_oracle(hb, oracle, world):
    if _n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    data = oracle.synth(9001, 67)
    q = oracle.synth(70, 68)
    multi = hb.MultiIndex(list(range(world)))
    for k in (10, 100):
        ek, ed = oracle.ground_truth(data, q, k)
        gk, gd = multi.ground_truth(data, q, k)
        assert np.array_equal(gk, ek) and np.array_equal(gd, ed), k
    ms = multi.phase_ms()
    assert len(ms) == 3 and ms[0] > 0


@pytest.mark.parametrize("world", [1, 2])
def test_build_sharded_recall(hb, oracle, world):
    if _n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    from oracle import oracle as orc
    data = oracle.synth(20000, 69)
    q = oracle.synth(200, 70)
    multi = hb.MultiIndex(list(range(world))).build_sharded(data, efc=64, m=12, M=48, serial_prefix=2000)
    gk, gd, gf = multi.knns_arrays(q, 10, 64)
    tk, td = oracle.ground_truth(data, q, 10)
    assert (gf == 10).all()
    # the same builder on one device over all rows is the yardstick (a sharded index searches every shard with the full
    # ef, so it can only find more)
    sk, _, _ = hb.build_hnsw(data, efc=64, m=12, M=48, serial_prefix=2000).knns_arrays(q, 10, 64)
    assert orc.recall_at_k(gk, tk, 10) >= orc.recall_at_k(sk, tk, 10) - 0.05
    # distances are recomputable from the global keys
    for i in range(0, 200, 37):
        for j in range(10):
            assert int(gd[i, j]) == oracle.hamming(data[int(gk[i, j])], q[i])


def test_concurrent_callers_on_one_index(hb, oracle):
    """Index::search is callable concurrently from any thread (src/index/mod.rs:21-43): per-call workspaces, no
    index-wide lock.  Eight host threads hammer one index with different batches; every answer equals the oracle's."""
    import threading
    data = oracle.synth(5000, 71)
    oidx = oracle.build(data, "hnsw", efc=32, m=8, M=24, parallel=False)
    idx = hb.index_from_parts("hnsw", oidx.graphs(), oidx.ep)
    batches = [oracle.synth(50 + 13 * t, 100 + t) for t in range(8)]
    want = [oidx.knns(b, 10, 64 if t % 2 else 300) for t, b in enumerate(batches)]   # both kernels (ef <= 256 and above)
    errors = []

    def work(t):
        try:
            for _ in range(5):
                gk, gd, gf = idx.knns_arrays(batches[t], 10, 64 if t % 2 else 300)
                assert np.array_equal(gk, want[t][0]) and np.array_equal(gd, want[t][1]) and np.array_equal(gf, want[t][2])
        except Exception as e:  # noqa: BLE001
            errors.append((t, repr(e)))

    th = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errors, errors


def test_library_calls_leave_the_callers_current_device_alone(hb, oracle):
    """Every entry point works on the device of its index / communicator and puts the caller's current device back: a
    host with its own notion of "current device" (PyTorch allocates on cudaGetDevice()) must not see it move."""
    if _n_devices() < 2:
        pytest.skip("needs 2 GPUs")
    import torch
    torch.cuda.set_device(1)
    data, q = oracle.synth(3000, 73), oracle.synth(40, 74)
    idx0 = hb.build_hnsw(data, efc=32, m=8, M=32, serial_prefix=500, device=0)
    assert torch.cuda.current_device() == 1
    idx0.knns_arrays(q, 10, 64)
    hb.ground_truth(data, q, 10, device=0)
    idx0.compact()
    assert torch.cuda.current_device() == 1
    t = torch.empty(4, device="cuda")
    assert t.device.index == 1
    del idx0
    assert torch.cuda.current_device() == 1
    torch.cuda.set_device(0)
