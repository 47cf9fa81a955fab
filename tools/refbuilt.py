"""configs[1] says "reference-built index": compare the graph hb_build_hnsw makes on the device with one made by the
restated reference builder (oracle HNSWBuilder::extend_parallel <- src/index/hnsw.rs:44-169) on the SAME data, both
searched by the CUDA path: parity vs the oracle on the reference-built .idx, QPS, X/E/V/A, recall, per-layer sizes and
base-degree deciles (cli/src/main.rs:769-786).  Runs on the GPU box.
usage: python tools/refbuilt.py N NQ > profiles/r02_refbuilt_<N>.json"""
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import hnsw_itu_b200 as hb  # noqa: E402
from oracle import oracle as orc  # noqa: E402

n, nq = int(sys.argv[1]), int(sys.argv[2])
k, ef = 10, 64
o = orc.Oracle(orc.build(march="native", out=os.path.join(os.path.dirname(orc.__file__), "liboracle_native.so")))
cores = len(os.sched_getaffinity(0))
d = torch.empty((n, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d.data_ptr(), n, 1003)
dq = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(dq.data_ptr(), nq, 1004)
torch.cuda.synchronize()
data, q = d.cpu().numpy().view(np.uint64), dq.cpu().numpy().view(np.uint64)
gk, _ = hb.ground_truth(data, q, k)


def deciles(deg):
    return [int(x) for x in np.percentile(deg, [0, 10, 20, 30, 40, 50, 60, 70, 80, 90, 100])]


def describe(idx):
    inf = idx.info()
    layers = []
    for g in range(inf.n_graphs):
        gr = idx.graph(g)
        deg = np.diff(gr["off"].astype(np.int64))
        layers.append({"graph": g, "nodes": int(len(deg)), "edges": int(deg.sum()), "mean_degree": float(deg.mean()) if len(deg) else 0.0,
                       "degree_deciles": deciles(deg) if len(deg) else []})
    return layers


def measure(idx):
    dk = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dd = torch.empty_like(dk)
    df = torch.empty(nq, dtype=torch.int32, device="cuda")
    dc = torch.empty((nq, 4), dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    idx.knns_device(dq.data_ptr(), nq, k, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), dc.data_ptr(), s)
    keys, dists = dk.cpu().numpy().view(np.uint64).copy(), dd.cpu().numpy().view(np.uint64).copy()
    c = dc.cpu().numpy().astype(np.float64).mean(0)
    for _ in range(3):
        idx.knns_device(dq.data_ptr(), nq, k, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        idx.knns_device(dq.data_ptr(), nq, k, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    return keys, dists, {"qps": nq / ms * 1e3, "ms_per_batch": ms, "recall_at_10": orc.recall_at_k(keys, gk, k),
                         "X": c[0], "E": c[1], "V_gpu": c[2], "A": c[3]}


out = {"n": n, "nq": nq, "k": k, "ef": ef, "params": "m=24 M=256 efc=96, chunk = 32 x 16 (rayon_threads = 16)", "host_cores": cores}
# ---- reference builder restatement on the host cores -> .idx -> hb_index_load_idx ------------------------------------
t0 = time.time()
oidx = o.build(data, "hnsw", 96, 24, 256, True, 16, cores)
t_ref = time.time() - t0
with tempfile.TemporaryDirectory() as tmp:
    path = os.path.join(tmp, "ref.idx")
    oidx.save_idx(path)
    out["idx_bytes"] = os.path.getsize(path)
    ridx = hb.read_index(path, keep_host=True)
rk, rd, rm = measure(ridx)
ok, od, of, oc = oidx.knns(q, k, ef, counters=True, threads=cores)
rm.update({"build_s": t_ref, "builder": f"oracle extend_parallel on {cores} host threads",
           "parity_vs_oracle": {"queries_identical_ids_and_distances": int(((ok == rk).all(1) & (od == rd).all(1)).sum()), "of": nq},
           "V_reference": float(oc[:, 2].mean()), "layers": describe(ridx)})
t0 = time.perf_counter()
oidx.knns(q, k, ef, threads=cores)
rm["cpu_oracle_qps"] = nq / (time.perf_counter() - t0)
out["reference_built"] = rm
del ridx
# ---- device builder on the same rows --------------------------------------------------------------------------------------
t0 = time.time()
gidx = hb.build_hnsw(data, efc=96, m=24, M=256, rayon_threads=16, keep_host=True)
t_gpu = time.time() - t0
_, _, gm = measure(gidx)
gm.update({"build_s": t_gpu, "builder": "hb_build_hnsw on the device", "layers": describe(gidx)})
goidx = o.from_parts("hnsw", [gidx.graph(g) for g in range(gidx.info().n_graphs)], gidx.info().entry_point)
_, _, _, goc = goidx.knns(q[:2000], k, ef, counters=True, threads=cores)
gm["V_reference"] = float(goc[:, 2].mean())
t0 = time.perf_counter()
goidx.knns(q, k, ef, threads=cores)
gm["cpu_oracle_qps"] = nq / (time.perf_counter() - t0)
out["device_built"] = gm
out["ratio_device_over_reference_built"] = {m_: gm[m_] / rm[m_] for m_ in ("qps", "X", "E", "V_gpu", "A", "V_reference", "recall_at_10", "cpu_oracle_qps")}
json.dump(out, sys.stdout, indent=1, default=float)
