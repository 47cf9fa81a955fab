"""Per-query start/end times from a -DHB_DIAG_TIME build (tools/, diagnostics): distribution of query latencies and
when the warp slots go idle.  HB_LIB_PATH=variants/time.so python tools/ab_time.py [EF]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hnsw_itu_b200 as hb
SHM = "/dev/shm/hb_ab"
ef = int(sys.argv[1]) if len(sys.argv) > 1 else 64
meta = np.load(f"{SHM}/meta.npy")
graphs = [{"sketches": np.load(f"{SHM}/g{g}_sk.npy", mmap_mode="r"), "down": np.load(f"{SHM}/g{g}_down.npy"),
           "off": np.load(f"{SHM}/g{g}_off.npy"), "ids": np.load(f"{SHM}/g{g}_ids.npy")} for g in range(int(meta[0]))]
q = np.load(f"{SHM}/queries.npy"); nq = len(q)
idx = hb.index_from_parts("hnsw", graphs, int(meta[1]), keep_host=False)
dq = torch.from_numpy(q.view(np.int64)).cuda()
dk = torch.empty((nq, 10), dtype=torch.int64, device="cuda"); dd = torch.empty_like(dk)
df = torch.empty(nq, dtype=torch.int32, device="cuda"); dc = torch.empty((nq, 4), dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), dc.data_ptr(), s)
c = dc.cpu().numpy().view(np.uint32).astype(np.int64)
t0 = c[:, 0].min()
start = ((c[:, 0] - t0) & 0xFFFFFFFF) / 1e3; end = ((c[:, 1] - t0) & 0xFFFFFFFF) / 1e3   # us
dur = end - start
print(f"kernel span {end.max():.0f} us; query latency us: mean {dur.mean():.0f} p10 {np.percentile(dur,10):.0f} p50 {np.percentile(dur,50):.0f} "
      f"p90 {np.percentile(dur,90):.0f} p99 {np.percentile(dur,99):.0f} max {dur.max():.0f}")
print("latency by start-time decile (us):", [f"{dur[(start >= a) & (start < b)].mean():.0f}" for a, b in zip(np.linspace(0, start.max() + 1, 11)[:-1], np.linspace(0, start.max() + 1, 11)[1:]) if ((start >= a) & (start < b)).any()])
slot = c[:, 2]; wib = slot % 4
print("latency by warp-in-block:", [f"{dur[wib == w].mean():.0f}" for w in range(4)], " by X corr:", np.corrcoef(dur, c[:, 3])[0, 1].round(3))
last = np.zeros(slot.max() + 1); np.maximum.at(last, slot, end)
print("slot idle-from (us) percentiles:", [f"{np.percentile(last, p):.0f}" for p in (0, 10, 50, 90, 100)], "queries/slot", np.bincount(slot).min(), np.bincount(slot).max())
act = [(np.sum((start <= t) & (end > t))) for t in np.linspace(0, end.max(), 21)]
print("active queries over time:", act)
