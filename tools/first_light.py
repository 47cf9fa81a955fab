"""Scratch timing run used early in round 1 (not part of the bench contract): builds an index with
the CPU oracle, moves it to the GPU, checks parity on a sample and times the search kernel."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hnsw_itu_b200 as hb
from oracle import oracle as orc

N = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
o = orc.get()
print("host threads", o.max_threads(), flush=True)
t = time.time(); data = o.synth(N, 1003); q = o.synth(nq, 1004); print("synth", time.time() - t, flush=True)
t = time.time(); oidx = o.build(data, "hnsw", 96, 24, 256, True, 16); print("cpu build", time.time() - t, flush=True)
t = time.time(); g = hb.index_from_parts("hnsw", oidx.graphs(), oidx.ep, keep_host=False); print("to gpu", time.time() - t, g.info().row_stride, g.info().device_bytes / 1e9, "GB", flush=True)
for ef in (16, 64, 128):
    gk, gd, gf, gc = g.knns_arrays(q, 10, ef, counters=True)
    t = time.time(); ok, od, of, oc = oidx.knns(q, 10, ef, counters=True); cpu_t = time.time() - t
    same = np.array_equal(gk, ok) and np.array_equal(gd, od) and np.array_equal(gc.astype(np.uint64)[:,[0,1,3]], oc[:,[0,1,3]])
    dq = torch.from_numpy(q.view(np.int64)).cuda()
    dk = torch.empty((nq, 10), dtype=torch.int64, device="cuda"); dd = torch.empty_like(dk); df = torch.empty(nq, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        g.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    c = gc.astype(np.float64).mean(0)
    bytes_q = 128 * c[2] + 4 * c[1] + 8 * c[0] + 128 + 160
    print(f"ef={ef} parity={same} gpu {ms:.3f} ms/batch {nq / ms * 1e3:.0f} QPS  cpu {nq / cpu_t:.0f} QPS  XEVA={c}  "
          f"bytes/q={bytes_q:.0f} achieved={bytes_q * nq / ms / 1e6:.1f} GB/s", flush=True)
