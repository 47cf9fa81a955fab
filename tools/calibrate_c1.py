"""BASELINE.json configs[0]: synthetic 100 k x 1024-bit sketches, 1 k queries, m=24 M=256 c=96 e=64 k=10 on the CPU path
(the reference CLI's `query`, cli/src/main.rs:550-571: build_index -> query_index).  The reference is Rust and cannot be
built here, so the CPU path is the oracle restatement (same algorithm, OpenMP over queries ~ rayon).  This is the
calibration record BASELINE.md section 4 (c1) asks for: the generator parameters are frozen where the oracle reaches
recall@10 in [0.93, 0.97] at ef = 64 (README.md:160-165 reports 0.93356 on LAION).  CPU only.
usage: python tools/calibrate_c1.py > profiles/r02_c1_calibration.json"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc  # noqa: E402

o = orc.Oracle(orc.build(march="native", out=os.path.join(os.path.dirname(orc.__file__), "liboracle_native.so")))
cores = len(os.sched_getaffinity(0))
n, nq, k, ef = 100_000, 1_000, 10, 64
sp = orc.SynthParams()
data, q = o.synth(n, 1003, sp), o.synth(nq, 1004, sp)
out = {"config": "BASELINE.json configs[0]: 100k x 1024-bit, 1k queries, m=24 M=256 c=96 e=64 k=10, CPU path (oracle restatement)",
       "generator": {"latent_dim": sp.latent_dim, "n_clusters": sp.n_clusters, "centre_scale": sp.centre_scale, "noise_scale": sp.noise_scale,
                     "distinct_planes": sp.distinct_planes, "seeds": {"planes": 1001, "centres": 1002, "data": 1003, "queries": 1004}},
       "host_cores": cores}
for label, parallel in (("extend_parallel (chunks of 32 x 16 after 50 000 sequential inserts, hnsw.rs:44-169)", True),
                        ("extend (every point sequentially: `index -S`)", False)):
    t0 = time.time()
    idx = o.build(data, "hnsw", 96, 24, 256, parallel, 16, cores)
    tb = time.time() - t0
    gk, _ = o.ground_truth(data, q, k, threads=cores)
    res = {}
    for e in (16, 32, 64, 96, 128, 256):
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter()
            ks, ds, fs, cs = idx.knns(q, k, e, counters=True, threads=cores)
            best = min(best, time.perf_counter() - t0)
        res[e] = {"recall_at_10": orc.recall_at_k(ks, gk, k), "qps": nq / best, "X": float(cs[:, 0].mean()), "E": float(cs[:, 1].mean()),
                  "V": float(cs[:, 2].mean()), "A": float(cs[:, 3].mean())}
    t0 = time.perf_counter()
    idx.knns(q, k, ef, threads=1)
    one = nq / (time.perf_counter() - t0)
    out[label] = {"build_s": tb, "graphs": [int(len(g["sketches"])) for g in idx.graphs()], "by_ef": res, "qps_1_thread_ef64": one}
json.dump(out, sys.stdout, indent=1)
