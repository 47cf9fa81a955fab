"""CPU, world_size 2 over gloo: the multi-GPU host logic (hnsw_itu_b200/sharded.py) -- shard plan, all-gather layout of
the per-shard top-k lists and the index broadcast -- checked against the oracle.  The oracle plays the GPU here (per-shard
exact top-k and the merge) because these tests run without a device; the -m gpu tests cover the same flow on CUDA."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hnsw_itu_b200 import sharded


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _pack(keys, dists):
    return ((dists.astype(np.uint64) << np.uint64(32)) | keys.astype(np.uint64)).view(np.int64)


def _worker(rank, world, port, n, nq, k, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as orc
        o = orc.get()
        data = o.synth(n, 1003)
        q = o.synth(nq, 1004)
        # ---- sharded ground truth: shard -> local exact top-k (oracle) -> all-gather -> merge == global ground truth
        start, ln = sharded.shard_bounds(n, world)[rank]
        lk, ld = o.ground_truth(data[start:start + ln], q, k)
        local = torch.from_numpy(_pack(lk + np.uint64(start), ld))
        lists = sharded.all_gather_lists(local)
        assert tuple(lists.shape) == (world, nq, k)
        assert torch.equal(lists[rank], local)
        merged = np.sort(lists.numpy().view(np.uint64).transpose(1, 0, 2).reshape(nq, world * k), axis=1)[:, :k]
        gk, gd = o.ground_truth(data, q, k)
        assert np.array_equal(merged & np.uint64(0xFFFFFFFF), gk)
        assert np.array_equal(merged >> np.uint64(32), gd)
        # ---- replicated index: rank 0's graphs arrive bit-identical on rank 1; each rank searches its query slice
        oidx = o.build(data, "hnsw", efc=32, m=8, M=32, parallel=False)
        graphs, ep, algo = sharded.broadcast_graphs(oidx.graphs() if rank == 0 else None, oidx.ep if rank == 0 else None, src=0, algo="hnsw")
        assert algo == "hnsw" and ep == oidx.ep and len(graphs) == oidx.num_graphs()
        for a, b in zip(graphs, oidx.graphs()):
            for key in ("sketches", "down", "off"):
                assert np.array_equal(np.asarray(a[key], dtype=np.uint64), np.asarray(b[key], dtype=np.uint64))
            assert np.array_equal(np.asarray(a["ids"], dtype=np.uint64), np.asarray(b["ids"], dtype=np.uint64))
        qs, ql = sharded.query_slice(nq, rank, world)
        rk, rd, rf = o.from_parts("hnsw", graphs, ep).knns(q[qs:qs + ql], k, 32)
        parts = [None] * world
        dist.all_gather_object(parts, (qs, rk, rd))
        full_k = np.concatenate([p[1] for p in sorted(parts, key=lambda p: p[0])])
        ek, ed, ef_ = oidx.knns(q, k, 32)
        assert np.array_equal(full_k, ek)
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 100, 10_000_001):
        for w in (1, 2, 3, 8):
            b = sharded.shard_bounds(n, w)
            assert len(b) == w and b[0][0] == 0 and sum(l for _, l in b) == n
            assert all(b[i][0] + b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(l for _, l in b) - min(l for _, l in b) <= 1
    with pytest.raises(ValueError):
        sharded.shard_bounds(10, 0)


def test_world1_is_a_passthrough():
    t = torch.arange(12, dtype=torch.int64).reshape(3, 4)
    assert torch.equal(sharded.all_gather_lists(t), t.unsqueeze(0))
    assert sharded.broadcast_graphs([{"x": 1}], 3) == ([{"x": 1}], 3, "hnsw")


@pytest.mark.timeout(300)
def test_world2_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 3000, 40, 10, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
