"""Phase clocks of the tcgen05 ground-truth kernel from a -DHB_GT_DIAG build (diagnostics, not the product):
   tools/ab_build.sh gtdiag "-DHB_GT_DIAG" && HB_LIB_PATH=variants/gtdiag.so python tools/gt_diag.py [N] [NQ] [K] [mma|mxf4]"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hnsw_itu_b200 as hb
from hnsw_itu_b200 import _lib
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
NQ = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 10
ALGO = sys.argv[4] if len(sys.argv) > 4 else "mma"   # "mma" (gt_tcgen05.cu) or "mxf4" (gt_mxf4.cu)
L = _lib.lib()
d = torch.empty((N, 16), dtype=torch.int64, device="cuda"); q = torch.empty((NQ, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d.data_ptr(), N, 1003); hb.synth_sketches(q.data_ptr(), NQ, 1004)
out = torch.empty((NQ, K), dtype=torch.int64, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(2):
    e0.record()
    _lib.check(L.hb_ground_truth_device_algo(d.data_ptr(), N, 0, q.data_ptr(), NQ, K, {"mma": 2, "mxf4": 3}[ALGO], out.data_ptr(), None))
    e1.record(); torch.cuda.synchronize()
print("ms", e0.elapsed_time(e1))
buf = np.zeros(256 * 16, dtype=np.int64)
assert getattr(L, "hb_debug_gt_diag" if ALGO == "mma" else "hb_debug_gt4_diag")(buf.ctypes.data_as(C.c_void_p)) == 0
g = buf.reshape(256, 16)[:148]
names = ["kernel", "mma:wait_acc_free", "mma:wait_stage_full", "exp:wait_stage_empty", "exp:fence", "epi:wait_acc_full", "epi:insert", "tiles", "exp:wait_A_free", "exp:wait_raw_full", "exp:stores"]
tiles = g[:, 7].astype(float)
print("tiles/CTA min/mean/max", tiles.min(), tiles.mean(), tiles.max())
for i, n in enumerate(names):
    if i == 7: continue
    v = g[:, i].astype(float)
    print(f"{n:22s} cycles/tile mean {np.mean(v / tiles):9.0f}  min {np.min(v / tiles):9.0f} max {np.max(v / tiles):9.0f}   share of kernel {np.mean(v / g[:, 0]):.3f}")

tr = buf[2400:2400 + 32 * 32].reshape(32, 32).astype(np.int64)
if tr[:, 0].any():
    t0 = tr[0, 0]
    print("trace of CTA 0 (cycles relative to the first traced tile): acc_free_pass | stage-full passes kb0..7 | last issue | epi sees full w0..3 | epi releases w0..3 | expansion arrives kb0..7")
    for i in range(0, 12):
        r = tr[i] - t0
        print(f"tile {i:2d}: A {r[0]:7d} | F {' '.join(f'{x:6d}' for x in r[18:26])} | I {r[1]:7d} | ef {' '.join(f'{x:6d}' for x in r[2:6])} | er {' '.join(f'{x:6d}' for x in r[6:10])} | xa {' '.join(f'{x:6d}' for x in r[10:18])}")
