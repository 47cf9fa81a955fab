"""End-to-end time of hb_ground_truth on HOST buffers (what the CLI's `ground-truth` calls): upload + search + download.
python tools/gt_host_bench.py [N] [NQ] [K]   (HB_LIB_PATH selects a variant build, e.g. -DHB_GT_SLICE_LOG2=40 = one slice)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hnsw_itu_b200 as hb
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
NQ = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 10
d = torch.empty((N, 16), dtype=torch.int64, device="cuda"); q = torch.empty((NQ, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d.data_ptr(), N, 1003); hb.synth_sketches(q.data_ptr(), NQ, 1004)
hd, hq = d.cpu().numpy().view(np.uint64), q.cpu().numpy().view(np.uint64)
del d, q
torch.cuda.empty_cache()
ts = []
for _ in range(4):
    t0 = time.perf_counter(); keys, dists = hb.ground_truth(hd, hq, K); ts.append((time.perf_counter() - t0) * 1e3)
import hashlib
print({"n": N, "nq": NQ, "k": K, "ms_all": [round(t, 1) for t in ts], "ms_best": round(min(ts[1:]), 1),
       "sha": hashlib.sha256(((dists << np.uint64(32)) | keys).tobytes()).hexdigest()[:16]})
