"""Scratch profiling driver (tools/, not the bench): oracle-built index -> GPU, a few ef=64 batches."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hnsw_itu_b200 as hb
from oracle import oracle as orc
N = int(sys.argv[1]); nq = int(sys.argv[2]); ef = int(sys.argv[3]); reps = int(sys.argv[4])
o = orc.get()
data = o.synth(N, 1003); q = o.synth(nq, 1004)
oidx = o.build(data, "hnsw", 96, 24, 256, True, 16)
g = hb.index_from_parts("hnsw", oidx.graphs(), oidx.ep, keep_host=False)
dq = torch.from_numpy(q.view(np.int64)).cuda()
dk = torch.empty((nq, 10), dtype=torch.int64, device="cuda"); dd = torch.empty_like(dk); df = torch.empty(nq, dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
for i in range(reps):
    e0.record()
    g.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e1.record(); torch.cuda.synchronize()
    print("ms", e0.elapsed_time(e1), flush=True)
