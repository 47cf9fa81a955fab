"""Ground-truth (brute-force top-k) timing on one GPU: the tcgen05 int8-GEMM kernel vs the XOR+POPC scan, with an
exact cross-check of the two result sets.  tools/, not the bench: python tools/gt_bench.py [N] [NQ] [K] [REPS] [algos]
Reports ms, int8 tensor op/s (2*nq*N*1024 per call) and the fraction of the nominal 4.5 Pop/s dense int8 peak."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hnsw_itu_b200 import _lib

N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
NQ = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 10
REPS = int(sys.argv[4]) if len(sys.argv) > 4 else 3
ALGOS = (sys.argv[5] if len(sys.argv) > 5 else "mma,scan").split(",")
CODE = {"auto": 0, "scan": 1, "mma": 2, "mxf4": 3}

import hnsw_itu_b200 as hb
L = _lib.lib()
d_data = torch.empty((N, 16), dtype=torch.int64, device="cuda")
d_q = torch.empty((NQ, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d_data.data_ptr(), N, 1003)
hb.synth_sketches(d_q.data_ptr(), NQ, 1004)
torch.cuda.synchronize()

out = {}
res = {}
for algo in ALGOS:
    packed = torch.empty((NQ, K), dtype=torch.int64, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for r in range(REPS + 1):
        e0.record()
        _lib.check(L.hb_ground_truth_device_algo(d_data.data_ptr(), N, 0, d_q.data_ptr(), NQ, K, CODE[algo], packed.data_ptr(), None))
        e1.record()
        torch.cuda.synchronize()
        if r:
            times.append(e0.elapsed_time(e1))
    ms = min(times)
    ops = 2.0 * NQ * N * 1024
    out[algo] = {"ms_best": ms, "ms_all": times, "pair_per_s": NQ * N / ms * 1e3, "int8_op_per_s": ops / ms * 1e3,
                 "frac_of_nominal_int8_4.5Pop": ops / ms * 1e3 / 4.5e15}
    res[algo] = packed.cpu().numpy()
if len(res) > 1:
    a, b = list(res)[:2]
    out["identical"] = bool(np.array_equal(res[a], res[b]))
print(json.dumps({"n": N, "nq": NQ, "k": K, **out}))
