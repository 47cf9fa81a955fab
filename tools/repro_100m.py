"""Reproduce the configs[4] leg of bench.py on ONE GPU (one rank of the 8): shard of 12.5 M rows at a given offset, optional
device build of the shard index first, sharded ground truth through hb_comm (one rank: no NCCL), sharded search.
usage: python tools/repro_100m.py OFFSET BUILD(0|1) [ROWS]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import hnsw_itu_b200 as hb  # noqa: E402

off, build = int(sys.argv[1]), int(sys.argv[2])
rows = int(sys.argv[3]) if len(sys.argv) > 3 else 12_500_000
nq, k = 10_000, 10
comm = hb.Comm.from_torch(0)
dq = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(dq.data_ptr(), nq, 1004)
d_shard = torch.empty((rows, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d_shard.data_ptr(), rows, 1003, start=off)
torch.cuda.synchronize()
print("synth ok", flush=True)
sidx = None
if build:
    t0 = time.time()
    sidx = hb.build_hnsw(d_shard.cpu().numpy().view(np.uint64), efc=96, m=24, M=256, rayon_threads=16, keep_host=False)
    torch.cuda.synchronize()
    print(f"build ok {time.time() - t0:.1f}s", flush=True)
d_gt = torch.empty((nq, k), dtype=torch.int64, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for i in range(2):
    comm.ground_truth_sharded(d_shard.data_ptr(), rows, off, dq.data_ptr(), nq, k, d_gt.data_ptr(), s)
    torch.cuda.synchronize()
    print("gt ok", i, comm.phase_ms(), flush=True)
gt = d_gt.cpu().numpy().view(np.uint64) & np.uint64(0xFFFFFFFF)
if sidx is not None:
    d_out = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    for i in range(2):
        comm.knns_sharded(sidx, off, dq.data_ptr(), nq, k, 64, d_out.data_ptr(), s)
        torch.cuda.synchronize()
        print("knns ok", i, comm.phase_ms(), flush=True)
    keys = d_out.cpu().numpy().view(np.uint64) & np.uint64(0xFFFFFFFF)
    print("recall", np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(keys, gt)]) / 10)
print("done")
