#!/usr/bin/env python
"""bench.py -- queries/sec of the batched HNSW query path (BASELINE.json metric) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path
  python bench.py --impl reference [...]                         the reference algorithm on the host cores (CPU oracle)

Workload (BASELINE.json configs[1]): synthetic 10 M x 1024-bit sketches of laion2B-hammingv2 shape, 10 k queries,
index built with m=24 M=256 efc=96, searched with ef=64 k=10.  A "step" is one Index::knns pass over the 10 k-query
batch.  `value` has index and queries resident in HBM; `e2e` goes through the C-ABI hb_knns with pinned HOST buffers
(H2D of the queries and D2H of ids/distances inside the timed region).  The index is built on the device by
hb_build_hnsw (the reference cannot be compiled here, so there is no Rust-built .idx to load); both arms search an
index built that way from the same data and parameters.  With N > 1 the index is replicated (built once per rank)
and every rank searches its own 10 k-query batch: no data-path collective, weak scaling.

Only the cpu_baseline / --impl reference legs touch oracle/ (CPU restatement of the reference), as the timed CPU
baseline and as the checker of bit-exact parity; the product path never does.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "queries/sec at recall@10>=0.93 (10Mx1024-bit, ef=64)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=10_000_000, help="data set size (default: the 10 M configuration)")
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--ef", type=int, default=64)
    ap.add_argument("--efc", type=int, default=96)
    ap.add_argument("--m", type=int, default=24)
    ap.add_argument("--M", type=int, default=256)
    ap.add_argument("--rayon-threads", type=int, default=16, help="stands in for rayon's thread count at build time (chunk = 32x)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def synth_on_device(hb, torch, n, seed):
    t = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(t.data_ptr(), n, seed)
    torch.cuda.synchronize()
    return t


def recall_at_k(ids, gt, k):
    """cli/eval/eval.py:48-56"""
    return float(np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(ids[:, :k], gt[:, :k])]) / k)


def native_oracle():
    """The CPU restatement compiled -O3 -march=native on THIS box (mirrors RUSTFLAGS=-Ctarget-cpu=native, README.md:81-84)."""
    from oracle import oracle as orc
    out = os.path.join(ROOT, "oracle", "liboracle_native.so")
    try:
        path = orc.build(march="native", out=out, force=not os.path.exists(out))
    except Exception:
        path = orc.build()
    return orc, orc.Oracle(path)


def oracle_index_from(idx, o):
    inf = idx.info()
    graphs = [idx.graph(g) for g in range(inf.n_graphs)]
    return o.from_parts("hnsw", graphs, None if inf.entry_point < 0 else inf.entry_point)


def build_index(hb, torch, a, keep_host):
    d = synth_on_device(hb, torch, a.n, 1003)
    data = d.cpu().numpy().view(np.uint64)
    t0 = time.time()
    idx = hb.build_hnsw(data, efc=a.efc, m=a.m, M=a.M, rayon_threads=a.rayon_threads, keep_host=keep_host)
    return d, data, idx, time.time() - t0


def workload_name(a):
    return (f"synthetic {a.n} x 1024-bit sketches (SimHash of a 4096-cluster latent mixture, seeds 1001-1004), {a.nq} queries per GPU, "
            f"index m={a.m} M={a.M} efc={a.efc} built on the device, ef={a.ef} k={a.k}")


def run_reference(a, rank, world):
    """The reference algorithm on the host cores: oracle knns (OpenMP dynamic over queries ~ rayon par_iter, mod.rs:36-41)."""
    if rank != 0:
        return
    import torch
    import hnsw_itu_b200 as hb
    torch.cuda.set_device(0)
    _, _, idx, _ = build_index(hb, torch, a, keep_host=True)
    q = synth_on_device(hb, torch, a.nq, 1004).cpu().numpy().view(np.uint64)
    orc, o = native_oracle()
    oidx = oracle_index_from(idx, o)
    del idx
    cores = o.max_threads()
    for _ in range(a.warmup):
        oidx.knns(q, a.k, a.ef)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        oidx.knns(q, a.k, a.ef)
    dt = time.perf_counter() - t0
    qps = a.nq * a.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 xor+popcount",
            "data": "synthetic", "config": {"workload": workload_name(a), "note": "CPU oracle (C++ restatement of the reference, -O3 -march=native, OpenMP) on rank 0 only"},
            "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port",
                             "sample": f"the full {a.nq}-query batch, {a.steps} passes"},
            "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "reference":
        return run_reference(a, rank, world)

    import torch
    import torch.distributed as dist
    import hnsw_itu_b200 as hb
    from hnsw_itu_b200 import _lib
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = _lib.lib()
    want_cpu = (not a.no_cpu_baseline) and world == 1 and rank == 0

    # ---- setup (untimed): data, index, queries, ground truth -------------------------------------------------
    # N > 1: rank 0 builds the index once and its arrays are broadcast over NCCL/NVLink (replicated index); every rank
    # then searches its own batch of nq queries with no data-path collective.
    from hnsw_itu_b200 import sharded
    dq = synth_on_device(hb, torch, a.nq, 1004 + rank)
    idx, build_s, gt_keys, data, gt_ms, gt_e2e_ms = None, 0.0, None, None, None, None
    if rank == 0:
        d_data, data, idx, build_s = build_index(hb, torch, a, keep_host=True)
        # the same ground truth end to end through hb_ground_truth on HOST buffers (what the CLI's `ground-truth`
        # calls): sliced upload overlapped with the search, results back on the host; wall clock, best of two after a warm-up
        gt_e2e_ms = None
        if world == 1:
            hq = dq.cpu().numpy().view(np.uint64)
            for i in range(3):
                t0 = time.perf_counter()
                ek, _ = hb.ground_truth(data, hq, a.k)
                dt = (time.perf_counter() - t0) * 1e3
                if i:
                    gt_e2e_ms = dt if gt_e2e_ms is None else min(gt_e2e_ms, dt)
        if not want_cpu:
            del data
        gt = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")
        # exact ground truth for the recall figure = BASELINE.json configs[3] on one GPU (the `ground-truth` path as a
        # tcgen05 GEMM on E2M1 operands with fused top-k); timed here with CUDA events, best of the three calls after one
        # warm-up call, reported beside the headline
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        gt_ms = None
        for i in range(4):
            g0.record()
            _lib.check(lib.hb_ground_truth_device(d_data.data_ptr(), a.n, 0, dq.data_ptr(), a.nq, a.k, gt.data_ptr(), None))
            g1.record()
            torch.cuda.synchronize()
            if i:
                gt_ms = g0.elapsed_time(g1) if gt_ms is None else min(gt_ms, g0.elapsed_time(g1))
        gt_keys = gt.cpu().numpy().view(np.uint64) & np.uint64(0xFFFFFFFF)
        del d_data, gt
        torch.cuda.empty_cache()
    t_b = time.time()
    idx = sharded.broadcast_index(idx, src=0)
    bcast_s = time.time() - t_b
    dk = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")
    dd = torch.empty_like(dk)
    df = torch.empty(a.nq, dtype=torch.int32, device="cuda")
    dc = torch.empty((a.nq, 4), dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    h, qp, kp, dp, fp = idx._h, dq.data_ptr(), dk.data_ptr(), dd.data_ptr(), df.data_ptr()

    def step():
        _lib.check(lib.hb_knns_device(h, qp, a.nq, a.k, a.ef, 0, kp, dp, fp, None, stream))

    _lib.check(lib.hb_knns_device(h, qp, a.nq, a.k, a.ef, 0, kp, dp, fp, dc.data_ptr(), stream))  # counted pass
    gpu_keys = dk.cpu().numpy().view(np.uint64).copy()
    gpu_dists = dd.cpu().numpy().view(np.uint64).copy()
    ctr = dc.cpu().numpy().astype(np.float64)
    recall = recall_at_k(gpu_keys, gt_keys, a.k) if rank == 0 else None

    # ---- device-resident timing -------------------------------------------------------------------------------
    for _ in range(max(a.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = lib.hb_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    e0.record()
    for _ in range(a.steps):
        step()
        kernel_ms.append(lib.hb_index_last_search_ms(h))
    e1.record()
    torch.cuda.synchronize()
    launches = lib.hb_kernel_launches() - launches0
    total_ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([total_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        dist.barrier()
    clocks = sampler.stop()

    # ---- end to end through the C ABI with pinned host buffers ----------------------------------------------
    hq = torch.empty((a.nq, 16), dtype=torch.int64).pin_memory()
    hq.copy_(dq.cpu())
    hk = torch.empty((a.nq, a.k), dtype=torch.int64).pin_memory()
    hd = torch.empty((a.nq, a.k), dtype=torch.int64).pin_memory()
    hf = torch.empty(a.nq, dtype=torch.int32).pin_memory()
    import ctypes as C
    u64p, u32p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
    args = (h, C.cast(hq.data_ptr(), u64p), a.nq, a.k, a.ef, 0, C.cast(hk.data_ptr(), u64p), C.cast(hd.data_ptr(), u64p),
            C.cast(hf.data_ptr(), u32p), None)
    for _ in range(max(a.warmup, 3)):
        _lib.check(lib.hb_knns(*args))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        _lib.check(lib.hb_knns(*args))
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    assert np.array_equal(hk.numpy().view(np.uint64), gpu_keys), "host-buffer path returned different ids"

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- CPU baseline + parity (rank 0, N = 1): the oracle on the same index, all host threads -----------------
    cpu, parity, v_alg = None, None, None
    if want_cpu:
        orc, o = native_oracle()
        oidx = oracle_index_from(idx, o)
        q = dq.cpu().numpy().view(np.uint64)
        ok, od, of, oc = oidx.knns(q, a.k, a.ef, counters=True)
        same = int(((ok == gpu_keys).all(axis=1) & (od == gpu_dists).all(axis=1)).sum())
        xea = bool(np.array_equal(oc[:, [0, 1, 3]], ctr[:, [0, 1, 3]].astype(np.uint64)))
        parity = {"queries_identical_ids_and_distances": same, "of": a.nq, "expansion_counters_identical": xea}
        v_alg = float(oc[:, 2].mean())
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            oidx.knns(q, a.k, a.ef)
            best = min(best, time.perf_counter() - t0)
        t0 = time.perf_counter()
        oidx.knns(q[:1000], a.k, a.ef, threads=1)
        t1 = time.perf_counter() - t0
        cpu = {"value": a.nq / best, "unit": "queries/s", "cores": o.max_threads(), "kind": "port",
               "sample": f"the full {a.nq}-query batch on the same index, best of 3; 1 thread: {1000 / t1:.0f} queries/s on 1000 queries"}

    # ---- roofline of the search kernel: algorithmic gather bytes / kernel time --------------------------------
    X, E, Vgpu = ctr[:, 0].mean(), ctr[:, 1].mean(), ctr[:, 2].mean()
    v_note = "oracle |visited| over the whole batch"
    if v_alg is None:
        # no full CPU leg (N > 1 or --no-cpu-baseline): the reference's |visited| from the oracle on a 500-query sample,
        # scaled by the GPU's own evaluation counts (the GPU re-evaluates nodes its recently-seen cache forgot)
        try:
            orc, o = native_oracle()
            oidx = oracle_index_from(idx, o)
            ns = min(500, a.nq)
            _, _, _, oc = oidx.knns(dq.cpu().numpy().view(np.uint64)[:ns], a.k, a.ef, counters=True)
            v_alg = float(oc[:, 2].mean()) * float(Vgpu / ctr[:ns, 2].mean())
            v_note = f"oracle |visited| on the first {ns} queries, scaled to the batch by the GPU's evaluation counts"
        except Exception as e:  # host graphs not kept: fall back to the GPU's (larger) count and say so
            v_note = f"GPU distance evaluations (oracle sample unavailable: {type(e).__name__}); overstates the algorithmic bytes"
    V = v_alg if v_alg is not None else Vgpu
    bytes_q = 128 * V + 4 * E + 8 * X + 128 + 16 * a.k
    k_ms = float(np.mean(kernel_ms))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = bytes_q * a.nq / (k_ms * 1e-3) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "search_kernel_ncu.json"))).get("dram_bytes_per_launch")
    except Exception:
        pass
    value = world * a.nq * a.steps / (total_ms * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": total_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64 xor+popcount (integer exact)", "data": "synthetic",
        "config": {"workload": workload_name(a), "recall_at_10": recall, "l2": "inputs larger than L2: the index is "
                   f"{idx.info().device_bytes / 1e9:.1f} GB and one step gathers {bytes_q * a.nq / 1e9:.2f} GB",
                   "index_build_s": build_s, "index_broadcast_s": bcast_s if world > 1 else None, "parallelism": "replicated index (built on rank 0, broadcast over NCCL), one query batch per GPU, no data-path collective" if world > 1 else "1 GPU",
                   "per_query": {"expansions": X, "ids_scanned": E, "visited_reference": V, "visited_reference_source": v_note,
                                 "distance_evals_gpu": Vgpu}},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": "hnsw_search_fast_kernel<2>", "kernel_ms": k_ms, "bytes_per_query": bytes_q,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "cpu_baseline": cpu,
        "e2e": {"value": world * a.nq * a.steps / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": a.nq * 128,
                "d2h_bytes_per_step": a.nq * a.k * 16 + a.nq * 4, "ms_per_step": e2e_s / a.steps * 1e3,
                "how": "hb_knns on pinned host buffers: queries are read from and results written to host memory across PCIe "
                       "inside the timed region (the search kernel accesses the mapped buffers in place; pageable buffers are staged with cudaMemcpyAsync)"},
        "gpu_launches": int(launches), "clocks": clocks, "parity": parity,
    }
    if gt_ms is not None:
        ops = 2.0 * a.nq * a.n * 1024
        tops = ops / (gt_ms * 1e-3) / 1e12
        bf16 = float(peaks.get("bf16_tflops", 1590.0))
        mma = a.nq >= 32 and a.n >= 4096
        line["ground_truth"] = {
            "workload": f"brute-force Hamming top-{a.k}, {a.n} x {a.nq} (BASELINE.json configs[3] on one GPU), untimed setup of the headline",
            "kernel": "gt_mxf4_kernel (tcgen05.mma kind::mxf4.block_scale on E2M1 operands with unit scales, TMEM accumulators, fused top-k)" if mma else "scan_topk_kernel",
            "ms": gt_ms, "pairs_per_s": a.nq * a.n / (gt_ms * 1e-3),
            "e2e_host_buffers_ms": gt_e2e_ms, "e2e_note": "hb_ground_truth on pageable host arrays: upload of the rows in slices by "
            "host threads through page-locked slots, overlapped with the search of the previous slice, results copied back; wall clock",
            "roofline": {"bound": "tensor", "achieved": tops, "unit": "Top/s on 4-bit operands (2*nq*N*1024 per call)",
                         "peak": 4.0 * bf16, "frac": tops / (4.0 * bf16),
                         "peak_source": "4 x MEASURED_PEAKS.json bf16_tflops (no fp4 figure is measured; nominal dense fp4 is 9000, int8 4500)",
                         "frac_of_nominal_9000": tops / 9000.0,
                         "note": "tensor pipe 80 % active in the ncu capture (profiles/r01k_gt_mxf4_kernel_ncu_summary.json); the rest is the "
                                 "epilogue at the TMEM read rate and the operand expansion sharing issue slots (DESIGN.md 4.3)"}}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
