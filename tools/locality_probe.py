"""Experiment (tools/, not the bench): does memory locality of the node order matter to the search kernel?
Build the 10 M index, take a breadth-first order of its base graph from the entry point, permute the DATA rows into that
order, rebuild, and time both indexes on the same queries.  The rebuilt graph is a different (comparable) graph, so the
figure of merit is achieved gather bandwidth (bytes from the kernel's own counters / time), not QPS alone.
usage (GPU box): python tools/locality_probe.py [N] [NQ]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import hnsw_itu_b200 as hb  # noqa: E402
from scipy.sparse import csr_matrix  # noqa: E402
from scipy.sparse.csgraph import breadth_first_order  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
d = torch.empty((n, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d.data_ptr(), n, 1003)
dq = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(dq.data_ptr(), nq, 1004)
torch.cuda.synchronize()
data = d.cpu().numpy().view(np.uint64)
q = dq.cpu().numpy().view(np.uint64)
gk, _ = hb.ground_truth(data, q, 10)


def measure(tag, idx, gt, rep):
    qq = dq.repeat(rep, 1) if rep > 1 else dq
    m = len(qq)
    dk = torch.empty((m, 10), dtype=torch.int64, device="cuda")
    dd = torch.empty_like(dk)
    df = torch.empty(m, dtype=torch.int32, device="cuda")
    dc = torch.empty((m, 4), dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    idx.knns_device(qq.data_ptr(), m, 10, 64, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), dc.data_ptr(), s)
    c = dc.cpu().numpy().astype(np.float64).mean(0)
    keys = dk.cpu().numpy().view(np.uint64)[:nq]
    for _ in range(3):
        idx.knns_device(qq.data_ptr(), m, 10, 64, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        idx.knns_device(qq.data_ptr(), m, 10, 64, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    bq = 128 * c[2] + 4 * c[1] + 8 * c[0] + 128 + 160
    rec = np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(keys, gt)]) / 10
    print(f"{tag:34s} nq={m:7d} {ms:8.3f} ms {m / ms / 1e3:7.3f} MQPS  X={c[0]:.1f} E={c[1]:.0f} Vgpu={c[2]:.0f}  gathered {bq * m / ms / 1e6:7.1f} GB/s  recall {rec:.4f}", flush=True)


t0 = time.time()
idx = hb.build_hnsw(data, efc=96, m=24, M=256, rayon_threads=16, keep_host=True)
print(f"build {time.time() - t0:.1f}s", flush=True)
measure("insertion order", idx, gk, 1)
measure("insertion order (saturating)", idx, gk, 10)
g = idx.graph(0)
off, ids = g["off"].astype(np.int64), g["ids"].astype(np.int32)
t0 = time.time()
A = csr_matrix((np.ones(len(ids), dtype=np.int8), ids, off), shape=(n, n))
order = breadth_first_order(A, 0, directed=True, return_predecessors=False)
rest = np.setdiff1d(np.arange(n, dtype=np.int64), order, assume_unique=False)
order = np.concatenate([order.astype(np.int64), rest])
print(f"bfs order {time.time() - t0:.1f}s, reached {n - len(rest)}/{n}", flush=True)
del idx, g, A, off, ids
data2 = np.ascontiguousarray(data[order])
inv = np.empty(n, dtype=np.int64)
inv[order] = np.arange(n)
gk2 = inv[gk.astype(np.int64)].astype(np.uint64)     # the same ground truth under the new labels
t0 = time.time()
idx2 = hb.build_hnsw(data2, efc=96, m=24, M=256, rayon_threads=16, keep_host=False)
print(f"build (BFS-ordered rows) {time.time() - t0:.1f}s", flush=True)
measure("BFS order", idx2, gk2, 1)
measure("BFS order (saturating)", idx2, gk2, 10)
idx2.compact()
measure("BFS order, compact rows", idx2, gk2, 1)
measure("BFS order, compact rows (sat.)", idx2, gk2, 10)
