"""Scratch scale run (tools/, not the bench): device synth -> GPU build -> ground truth -> ef sweep with recall and QPS."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hnsw_itu_b200 as hb

N = int(sys.argv[1]); nq = int(sys.argv[2]); threads = int(sys.argv[3]) if len(sys.argv) > 3 else 16
prefix = int(sys.argv[4]) if len(sys.argv) > 4 else 0
efs = [int(x) for x in sys.argv[5].split(",")] if len(sys.argv) > 5 else [16, 32, 64, 128]
t = time.time()
d = torch.empty((N, 16), dtype=torch.int64, device="cuda"); hb.synth_sketches(d.data_ptr(), N, 1003)
dq = torch.empty((nq, 16), dtype=torch.int64, device="cuda"); hb.synth_sketches(dq.data_ptr(), nq, 1004)
torch.cuda.synchronize()
data = d.cpu().numpy().view(np.uint64); q = dq.cpu().numpy().view(np.uint64)
print(f"synth {time.time()-t:.2f}s", flush=True)
t = time.time()
idx = hb.build_hnsw(data, 96, 24, 256, rayon_threads=threads, keep_host=False, serial_prefix=prefix)
bt = time.time() - t
inf = idx.info()
print(f"build {bt:.2f}s  graphs={inf.n_graphs} nodes={inf.n_nodes_total} edges={inf.n_edges_total} avgdeg={inf.n_edges_total/inf.n_nodes_total:.1f} "
      f"prunes={inf.build_prunes} unlinked={inf.build_unlinked} dropped={inf.build_dropped_appends} stride={inf.row_stride} dev={inf.device_bytes/1e9:.2f}GB", flush=True)
t = time.time()
lib = hb._lib.lib()
gt = torch.empty((nq, 10), dtype=torch.int64, device="cuda")
hb._lib.check(lib.hb_ground_truth_device(d.data_ptr(), N, 0, dq.data_ptr(), nq, 10, gt.data_ptr(), None)); torch.cuda.synchronize()
print(f"ground truth {time.time()-t:.2f}s", flush=True)
gtk = (gt.cpu().numpy().view(np.uint64) & np.uint64(0xFFFFFFFF))
dk = torch.empty((nq, 10), dtype=torch.int64, device="cuda"); dd = torch.empty_like(dk); df = torch.empty(nq, dtype=torch.int32, device="cuda")
dc = torch.empty((nq, 4), dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for ef in efs:
    idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), dc.data_ptr(), s)
    keys = dk.cpu().numpy().view(np.uint64)
    rec = np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(keys, gtk)]) / 10
    c = dc.cpu().numpy().astype(np.float64).mean(0)
    for _ in range(2):
        idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    bq = 128 * c[2] + 4 * c[1] + 8 * c[0] + 128 + 160
    print(f"ef={ef} recall@10={rec:.4f} {ms:.3f} ms/batch {nq/ms*1e3:.0f} QPS XEVA={c} bytes/q={bq:.0f} achieved={bq*nq/ms/1e6:.0f} GB/s", flush=True)
