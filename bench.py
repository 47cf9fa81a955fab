#!/usr/bin/env python
"""bench.py -- queries/sec of the batched HNSW query path (BASELINE.json metric) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path
  python bench.py --impl reference [...]                         the reference algorithm on the host cores (CPU oracle)

Workload (BASELINE.json configs[1]): synthetic 10 M x 1024-bit sketches of laion2B-hammingv2 shape, 10 k queries,
index built with m=24 M=256 efc=96, searched with ef=64 k=10.  A "step" is one Index::knns pass over a 10 k-query
batch.  `value` has index and queries resident in HBM; `e2e` goes through the C-ABI hb_knns with HOST buffers
(pageable, what a Rust Vec<Sketch> is; the pinned figure is reported beside it), H2D of the queries and D2H of
ids/distances inside the timed region.

N > 1 (one process per GPU, launched by torchrun): rank 0 builds the index, the library replicates it GPU-to-GPU
(hb_index_broadcast = ncclBroadcast inside libhnswb200), then
  * `value`                 weak scaling: every rank searches its own 10 k batch, no data-path collective;
  * `strong`                configs[2]: the ONE 10 k batch split N ways (contiguous slices, input order kept,
                            src/index/mod.rs:36-41), the saturating 1 M batch split N ways, and the ef sweep 16..256
                            (recall and QPS per ef) at that N;
  * `ground_truth_sharded`  configs[3]: rows sharded N ways, hb_ground_truth_sharded_device (tcgen05 kernel ->
                            ncclAllGather -> exact merge, all inside the library), ms + phase breakdown + a checksum
                            that must be the same at every N;
  * `sharded_hnsw_100M`     configs[4], N = 8 only: 8 shards x 12.5 M rows, per-shard device build, sharded search +
                            all-gather + merge, recall against the sharded exact ground truth.

Only the cpu_baseline / --impl reference legs touch oracle/ (CPU restatement of the reference), as the timed CPU
baseline and as the checker of bit-exact parity; the product path never does.
"""
from __future__ import annotations

import argparse
import ctypes as C
import hashlib
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
EF_SWEEP = [16, 24, 32, 48, 64, 96, 128, 192, 256]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--rows", dest="n", type=int, default=10_000_000,
                    help="data set size (default: the 10 M configuration); spell it --rows under torchrun, whose own parser rejects --n as ambiguous")
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--ef", type=int, default=64)
    ap.add_argument("--efc", type=int, default=96)
    ap.add_argument("--m", type=int, default=24)
    ap.add_argument("--M", type=int, default=256)
    ap.add_argument("--rayon-threads", type=int, default=16, help="stands in for rayon's thread count at build time (chunk = 32x)")
    ap.add_argument("--index-builder", default="device", choices=["device", "reference"],
                    help="device: hb_build_hnsw (38 s at 10 M); reference: the restated reference builder on the host cores, written as "
                         ".idx and loaded through hb_index_load_idx (~4 min at 10 M on 16 cores; profiles/r02_refbuilt_10M.json compares the two)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the strong / sweep / sharded legs (headline only)")
    ap.add_argument("--no-100m", action="store_true", help="skip configs[4] (100 M sharded HNSW) at N = 8")
    ap.add_argument("--shard-rows", type=int, default=12_500_000, help="rows per shard of configs[4]")
    ap.add_argument("--force-sharded-hnsw", action="store_true", help="run the configs[4] leg at any N (default: N = 8 only)")
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


def synth_on_device(hb, torch, n, seed, start=0):
    t = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(t.data_ptr(), n, seed, start=start)
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


def host_threads() -> int:
    """All logical CPUs this process may run on (rayon's default pool size); never OMP_NUM_THREADS, which torchrun sets to 1."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_index_from(idx, o):
    inf = idx.info()
    graphs = [idx.graph(g) for g in range(inf.n_graphs)]
    return o.from_parts("hnsw", graphs, None if inf.entry_point < 0 else inf.entry_point)


def build_index(hb, torch, a, keep_host):
    d = synth_on_device(hb, torch, a.n, 1003)
    data = d.cpu().numpy().view(np.uint64)
    t0 = time.time()
    if a.index_builder == "reference":
        # configs[1] "reference-built index": HNSWBuilder::extend_parallel restated (oracle), bincode .idx, read_index on the device
        import tempfile
        orc, o = native_oracle()
        oidx = o.build(data, "hnsw", a.efc, a.m, a.M, True, a.rayon_threads, host_threads())
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "reference_built.idx")
            oidx.save_idx(path)
            del oidx
            idx = hb.read_index(path, device=torch.cuda.current_device(), keep_host=keep_host)
    else:
        idx = hb.build_hnsw(data, efc=a.efc, m=a.m, M=a.M, rayon_threads=a.rayon_threads, keep_host=keep_host,
                            device=torch.cuda.current_device())
    return d, data, idx, time.time() - t0


def workload_name(a):
    return (f"synthetic {a.n} x 1024-bit sketches (SimHash of a 4096-cluster latent mixture, seeds 1001-1004), {a.nq} queries per GPU, "
            f"index m={a.m} M={a.M} efc={a.efc} built " + ("on the device" if a.index_builder == "device" else "by the restated reference builder on the host, loaded from its .idx") + f", ef={a.ef} k={a.k}")


def run_reference(a, rank, world):
    """The reference algorithm on the host cores: oracle knns (OpenMP dynamic over queries ~ rayon par_iter, mod.rs:36-41)
    with every logical CPU of the box, whatever OMP_NUM_THREADS says.  Rank 0 only; the index it searches is the one the
    `ours` arm searches (built on the device by hb_build_hnsw from the same data and parameters -- the Rust builder cannot
    run here; profiles/r02*_refbuilt*.json compares that graph with one from the restated CPU builder)."""
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
    cores = host_threads()
    for _ in range(a.warmup):
        oidx.knns(q, a.k, a.ef, threads=cores)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        oidx.knns(q, a.k, a.ef, threads=cores)
    dt = time.perf_counter() - t0
    qps = a.nq * a.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 xor+popcount",
            "data": "synthetic", "config": {"workload": workload_name(a), "note": f"CPU oracle (C++ restatement of the reference, -O3 -march=native, OpenMP, {cores} threads) on rank 0 only"},
            "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port",
                             "sample": f"the full {a.nq}-query batch, {a.steps} passes"},
            "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class Rig:
    """Per-rank state of the `ours` arm: torch, the library, the communicator, and device-timing helpers."""

    def __init__(self, a):
        import torch
        import torch.distributed as dist
        import hnsw_itu_b200 as hb
        from hnsw_itu_b200 import _lib
        self.a, self.torch, self.dist, self.hb, self._lib = a, torch, dist, hb, _lib
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        self.lib = _lib.lib()
        self.comm = hb.Comm.from_torch(self.local_rank)   # the library's own NCCL communicator (collectives run inside it)
        self.stream = torch.cuda.current_stream().cuda_stream

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps: int, warmup: int = 3) -> float:
        """ms per step: `warmup` untimed calls, then `steps` calls bracketed by barrier + synchronize, CUDA events on the
        launching stream, max over ranks."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = self.max_over_ranks(e0.elapsed_time(e1))
        self.barrier()
        return ms / steps

    def knns_fn(self, h, dq, nq, k, ef, dk, dd, df, counters=None, q_off=0):
        """A closure launching hb_knns_device on rows [q_off, q_off + nq) of dq into the same rows of dk/dd/df."""
        lib, check, s = self.lib, self._lib.check, self.stream
        qp, kp, dp, fp = dq.data_ptr() + q_off * 128, dk.data_ptr() + q_off * k * 8, dd.data_ptr() + q_off * k * 8, df.data_ptr() + q_off * 4
        cp = counters.data_ptr() + q_off * 16 if counters is not None else None

        def fn():
            check(lib.hb_knns_device(h, qp, nq, k, ef, 0, kp, dp, fp, cp, s))
        return fn

    def gather_rows(self, t, q_off, nq):
        """Every rank filled rows [q_off, q_off + nq) of its copy of t: sum the zero-padded copies into the full result."""
        if self.world == 1:
            return t
        full = self.torch.zeros_like(t)
        full[q_off:q_off + nq] = t[q_off:q_off + nq]
        self.dist.all_reduce(full, op=self.dist.ReduceOp.SUM)
        return full


_T0 = time.time()


def phase(msg):
    """Progress on stderr (rank 0): where the wall clock of a run goes."""
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench {time.time() - _T0:7.1f}s] {msg}", file=sys.stderr, flush=True)


def split(n, world, r):
    """Contiguous (start, len): the reference's `-b start -l len` (cli/src/main.rs:136-140)."""
    e0, e1 = (r * n) // world, ((r + 1) * n) // world
    return e0, e1 - e0


def main():
    a = parse()
    # NCCL writes its version banner (NCCL_DEBUG=VERSION/INFO) to stdout by default: keep stdout for the one JSON line
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if a.impl == "reference":
        return run_reference(a, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
    R = Rig(a)
    torch, hb, lib, _lib = R.torch, R.hb, R.lib, R._lib
    rank, world = R.rank, R.world
    want_cpu = (not a.no_cpu_baseline) and world == 1 and rank == 0
    u64 = np.uint64

    # ---- setup (untimed): index on rank 0, replicated GPU-to-GPU by the library ---------------------------------
    idx, build_s, data, gt_e2e_ms = None, 0.0, None, None
    dq = synth_on_device(hb, torch, a.nq, 1004)                       # THE 10 k batch (configs[1]); same on every rank
    dq_own = dq if rank == 0 else synth_on_device(hb, torch, a.nq, 1004 + rank)   # weak scaling: one batch per GPU
    if rank == 0:
        d_data, data, idx, build_s = build_index(hb, torch, a, keep_host=True)
        if world == 1:
            # `ground-truth` end to end through hb_ground_truth on pageable HOST arrays (what the CLI calls): sliced upload
            # overlapped with the search, results back on the host; wall clock, best of two after a warm-up
            hq = dq.cpu().numpy().view(u64)
            for i in range(3):
                t0 = time.perf_counter()
                hb.ground_truth(data, hq, a.k)
                dt = (time.perf_counter() - t0) * 1e3
                if i:
                    gt_e2e_ms = dt if gt_e2e_ms is None else min(gt_e2e_ms, dt)
        if not want_cpu:
            data = None
        del d_data
        torch.cuda.empty_cache()
    phase(f"index built in {build_s:.1f}s")
    t_b = time.time()
    idx = R.comm.broadcast_index(idx, 0)     # hb_index_broadcast: ncclBroadcast of sketches / adjacency rows / down-links
    R.barrier()
    bcast_s = time.time() - t_b
    h = idx._h

    phase(f"index replicated in {bcast_s:.1f}s")
    # ---- configs[3]: exact ground truth with the rows sharded over the ranks (also the recall reference) -------
    r0, rl = split(a.n, world, rank)
    d_shard = synth_on_device(hb, torch, rl, 1003, start=r0)
    d_gt = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")

    def gt_fn():
        R.comm.ground_truth_sharded(d_shard.data_ptr(), rl, r0, dq.data_ptr(), a.nq, a.k, d_gt.data_ptr(), R.stream)
    gt_ms = R.timed(gt_fn, 3, warmup=1)
    gt_phase = R.comm.phase_ms()
    gt_packed = d_gt.cpu().numpy().view(u64)
    gt_keys = gt_packed & u64(0xFFFFFFFF)
    gt_sha = hashlib.sha256(gt_packed.tobytes()).hexdigest()[:16]
    del d_shard
    torch.cuda.empty_cache()

    dk = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")
    dd = torch.empty_like(dk)
    df = torch.empty(a.nq, dtype=torch.int32, device="cuda")
    dc = torch.empty((a.nq, 4), dtype=torch.int32, device="cuda")

    # counted pass over the common batch: results for parity / recall, counters for the byte model
    R.knns_fn(h, dq, a.nq, a.k, a.ef, dk, dd, df, counters=dc)()
    torch.cuda.synchronize()
    gpu_keys = dk.cpu().numpy().view(u64).copy()
    gpu_dists = dd.cpu().numpy().view(u64).copy()
    ctr = dc.cpu().numpy().astype(np.float64)
    recall = recall_at_k(gpu_keys, gt_keys, a.k)
    ns = min(1000, a.nq)
    dce = torch.empty((ns, 4), dtype=torch.int32, device="cuda")
    _lib.check(lib.hb_knns_device(h, dq.data_ptr(), ns, a.k, a.ef, _lib.HB_EXACT_VISITED, dk.data_ptr(), dd.data_ptr(), df.data_ptr(),
                                  dce.data_ptr(), R.stream))
    assert np.array_equal(dk[:ns].cpu().numpy().view(u64), gpu_keys[:ns]), "exact-visited kernel returned different ids"
    v_exact = float(dce.cpu().numpy()[:, 2].mean()) * float(ctr[:, 2].mean() / ctr[:ns, 2].mean())

    phase("ground truth + counted pass done")
    # ---- headline: device-resident, one 10 k batch per GPU ------------------------------------------------------
    step = R.knns_fn(h, dq_own, a.nq, a.k, a.ef, dk, dd, df)
    warm = max(a.warmup, 3)
    for _ in range(warm):
        step()
    sampler = ClockSampler(R.local_rank)
    sampler.start()
    R.barrier()
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
    total_ms = R.max_over_ranks(e0.elapsed_time(e1))
    R.barrier()

    # ---- end to end through the C ABI with HOST buffers: pageable (the headline e2e) and pinned --------------------
    u64p, u32p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)

    def e2e_leg(hq, hk, hd, hf):
        args = (h, C.cast(hq.data_ptr(), u64p), a.nq, a.k, a.ef, 0, C.cast(hk.data_ptr(), u64p), C.cast(hd.data_ptr(), u64p),
                C.cast(hf.data_ptr(), u32p), None)
        for _ in range(warm):
            _lib.check(lib.hb_knns(*args))
        R.barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            _lib.check(lib.hb_knns(*args))
        torch.cuda.synchronize()
        return R.max_over_ranks(time.perf_counter() - t0)

    host = dict(hq=dq_own.cpu(), hk=torch.empty((a.nq, a.k), dtype=torch.int64), hd=torch.empty((a.nq, a.k), dtype=torch.int64),
                hf=torch.empty(a.nq, dtype=torch.int32))
    e2e_pageable_s = e2e_leg(**host)
    if rank == 0:
        assert np.array_equal(host["hk"].numpy().view(u64), gpu_keys), "pageable host-buffer path returned different ids"
    pinned = {k_: v.pin_memory() for k_, v in host.items()}
    e2e_pinned_s = e2e_leg(**pinned)
    if rank == 0:
        assert np.array_equal(pinned["hk"].numpy().view(u64), gpu_keys), "pinned host-buffer path returned different ids"

    phase("headline + e2e legs done")
    extras = {}
    if not a.no_extras:
        extras = run_extras(R, a, idx, dq, gpu_keys, gt_keys)
    # nvidia-smi samples every 100 ms and the timed region is K x 1.2 ms: the sampler runs from the start of the timed
    # region through the end-to-end and strong-scaling legs (all search load on the same GPU) to get a meaningful median
    clocks = sampler.stop()
    clocks["window"] = "from the start of the timed region to the end of the e2e / strong legs"

    if rank != 0:
        if world > 1:
            R.dist.destroy_process_group()
        return
    report(R, a, idx, dict(build_s=build_s, bcast_s=bcast_s, data=data, gt_e2e_ms=gt_e2e_ms, gt_ms=gt_ms, gt_phase=gt_phase,
                           gt_sha=gt_sha, gt_keys=gt_keys, v_exact=v_exact, gpu_keys=gpu_keys, gpu_dists=gpu_dists, ctr=ctr, recall=recall, kernel_ms=kernel_ms,
                           total_ms=total_ms, launches=launches, clocks=clocks, e2e_pageable_s=e2e_pageable_s,
                           e2e_pinned_s=e2e_pinned_s, warm=warm, want_cpu=want_cpu, dq=dq, extras=extras))
    if world > 1:
        R.dist.destroy_process_group()


def run_extras(R, a, idx, dq, gpu_keys, gt_keys):
    """configs[2] (split batch, saturating batch, ef sweep) and, at N = 8, configs[4] (100 M sharded HNSW)."""
    torch, hb, world, rank = R.torch, R.hb, R.world, R.rank
    u64 = np.uint64
    h = idx._h
    out = {}
    # ---- the ONE 10 k batch split N ways: contiguous slices, outputs in input order ---------------------------
    qs, ql = split(a.nq, world, rank)
    dk = torch.zeros((a.nq, a.k), dtype=torch.int64, device="cuda")
    dd = torch.zeros_like(dk)
    df = torch.zeros(a.nq, dtype=torch.int32, device="cuda")
    fn = R.knns_fn(h, dq, ql, a.k, a.ef, dk, dd, df, q_off=qs)
    ms = R.timed(fn, a.steps)
    same = bool(np.array_equal(R.gather_rows(dk, qs, ql).cpu().numpy().view(u64), gpu_keys))
    out["split_10k"] = {"queries": a.nq, "queries_per_gpu": ql, "ms_per_batch": ms, "qps": a.nq / ms * 1e3,
                        "identical_to_unsplit_ids": same,
                        "note": "one launch per GPU over its contiguous slice; below one wave per GPU the limiter is the per-query "
                                "dependent chain, not a collective (there is none)"}
    # ---- two concurrent callers per GPU, each with its own 10 k batch on its own stream (Index::search is Sync, mod.rs:21-43;
    #      per-call workspaces in the library): the second batch's head fills the first one's tail ----------------------
    import threading
    n_callers = 2
    streams = [torch.cuda.Stream() for _ in range(n_callers)]
    bufs = [(torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda"), torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda"),
             torch.empty(a.nq, dtype=torch.int32, device="cuda")) for _ in range(n_callers)]
    torch.cuda.synchronize()

    def caller(i, reps):
        k_, d_, f_ = bufs[i]
        for _ in range(reps):
            R._lib.check(R.lib.hb_knns_device(h, dq.data_ptr(), a.nq, a.k, a.ef, 0, k_.data_ptr(), d_.data_ptr(), f_.data_ptr(), None,
                                              streams[i].cuda_stream))

    def concurrent(reps):
        th = [threading.Thread(target=caller, args=(i, reps)) for i in range(n_callers)]
        [t.start() for t in th]
        [t.join() for t in th]
        torch.cuda.synchronize()
    concurrent(3)
    R.barrier()
    t0 = time.perf_counter()
    concurrent(a.steps)
    conc_s = R.max_over_ranks(time.perf_counter() - t0)
    out["concurrent_callers"] = {"callers_per_gpu": n_callers, "batches_per_caller": a.steps, "qps": world * n_callers * a.steps * a.nq / conc_s,
                                 "identical_ids": bool(np.array_equal(bufs[1][0].cpu().numpy().view(u64), gpu_keys)),
                                 "note": "wall clock over both callers; each call is still one launch + one stream synchronise"}
    # ---- saturating batch: the 10 k queries tiled x100 = 1 M, split N ways ------------------------------------
    tile = 100
    nq_sat = a.nq * tile
    ss, sl = split(nq_sat, world, rank)
    dq_sat = dq.repeat(tile, 1)[ss:ss + sl].contiguous()
    sk = torch.empty((sl, a.k), dtype=torch.int64, device="cuda")
    sd = torch.empty_like(sk)
    sf = torch.empty(sl, dtype=torch.int32, device="cuda")
    ms = R.timed(R.knns_fn(h, dq_sat, sl, a.k, a.ef, sk, sd, sf), 3, warmup=1)
    out["saturating_1M"] = {"queries": nq_sat, "queries_per_gpu": sl, "ms_per_batch": ms, "qps": nq_sat / ms * 1e3}
    # ---- ef sweep: recall and QPS per ef on the split 10 k batch and on the saturating batch -------------------
    sweep = []
    for ef in EF_SWEEP:
        ms10 = R.timed(R.knns_fn(h, dq, ql, a.k, ef, dk, dd, df, q_off=qs), 3, warmup=1)
        keys = R.gather_rows(dk, qs, ql).cpu().numpy().view(u64)
        mssat = R.timed(R.knns_fn(h, dq_sat, sl, a.k, ef, sk, sd, sf), 2, warmup=1)
        sweep.append({"ef": ef, "recall_at_10": recall_at_k(keys, gt_keys, a.k), "qps_split_10k": a.nq / ms10 * 1e3,
                      "qps_saturating_1M": nq_sat / mssat * 1e3})
    out["ef_sweep"] = sweep
    del dq_sat, sk, sd, sf
    torch.cuda.empty_cache()

    phase("split / concurrent / saturating / ef sweep done")
    # ---- configs[4]: 100 M rows as N shards, one HNSW per shard, sharded search + all-gather + exact merge ------
    if (world == 8 or a.force_sharded_hnsw) and not a.no_100m:
        rows = a.shard_rows
        off = rank * rows
        d_shard = synth_on_device(hb, torch, rows, 1003, start=off)
        t0 = time.time()
        sidx = hb.build_hnsw(d_shard.cpu().numpy().view(u64), efc=a.efc, m=a.m, M=a.M, rayon_threads=a.rayon_threads, keep_host=False,
                             device=R.local_rank)
        R.barrier()
        build_s = time.time() - t0
        d_gt = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")
        gt_fn = lambda: R.comm.ground_truth_sharded(d_shard.data_ptr(), rows, off, dq.data_ptr(), a.nq, a.k, d_gt.data_ptr(), R.stream)  # noqa: E731
        gt_ms = R.timed(gt_fn, 2, warmup=1)
        gt_phase = R.comm.phase_ms()
        gt100 = d_gt.cpu().numpy().view(u64) & u64(0xFFFFFFFF)
        d_out = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")
        fn = lambda: R.comm.knns_sharded(sidx, off, dq.data_ptr(), a.nq, a.k, a.ef, d_out.data_ptr(), R.stream)  # noqa: E731
        ms = R.timed(fn, 5, warmup=2)
        ph = R.comm.phase_ms()
        keys = d_out.cpu().numpy().view(u64) & u64(0xFFFFFFFF)
        out["sharded_hnsw_100M"] = {
            "rows": rows * world, "shards": world, "per_shard_build_s": build_s, "ms_per_batch": ms, "qps": a.nq / ms * 1e3,
            "recall_at_10": recall_at_k(keys, gt100, a.k), "ef": a.ef,
            "phase_ms_rank0": {"search": ph[0], "all_gather": ph[1], "merge": ph[2]},
            "ground_truth_100M": {"ms": gt_ms, "phase_ms_rank0": {"search": gt_phase[0], "all_gather": gt_phase[1], "merge": gt_phase[2]},
                                  "pairs_per_s": a.nq * rows * world / (gt_ms * 1e-3)},
            "note": "every rank searches all queries on its shard (the reference's `index -b -l` sub-range), local keys + offset, "
                    "ncclAllGather of nq*k packed u64 per rank + exact (distance, key) merge inside libhnswb200"}
        del sidx, d_shard
        torch.cuda.empty_cache()
        phase("sharded HNSW leg done")
    return out


def report(R, a, idx, s):
    """Rank 0: CPU baseline + parity (N = 1), roofline, the JSON line."""
    world = R.world
    u64 = np.uint64
    ctr, gpu_keys, gpu_dists = s["ctr"], s["gpu_keys"], s["gpu_dists"]
    cores = host_threads()

    # ---- CPU baseline + parity (rank 0, N = 1): the oracle on the same index, all host threads -----------------
    cpu, parity, v_alg, gt_cpu = None, None, None, None
    q = s["dq"].cpu().numpy().view(u64)
    if s["want_cpu"]:
        orc, o = native_oracle()
        oidx = oracle_index_from(idx, o)
        ok, od, of, oc = oidx.knns(q, a.k, a.ef, counters=True, threads=cores)
        same = int(((ok == gpu_keys).all(axis=1) & (od == gpu_dists).all(axis=1)).sum())
        xea = bool(np.array_equal(oc[:, [0, 1, 3]], ctr[:, [0, 1, 3]].astype(u64)))
        parity = {"queries_identical_ids_and_distances": same, "of": a.nq, "expansion_counters_identical": xea}
        v_alg = float(oc[:, 2].mean())
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            oidx.knns(q, a.k, a.ef, threads=cores)
            best = min(best, time.perf_counter() - t0)
        t0 = time.perf_counter()
        oidx.knns(q[:1000], a.k, a.ef, threads=1)
        t1 = time.perf_counter() - t0
        cpu = {"value": a.nq / best, "unit": "queries/s", "cores": cores, "kind": "port",
               "sample": f"the full {a.nq}-query batch on the same index, best of 3; 1 thread: {1000 / t1:.0f} queries/s on 1000 queries"}
        # ground truth on the CPU: oracle Bruteforce (per-query full scan + min_k, bruteforce.rs:37-47) on a bounded sample
        ns = min(1000, a.nq)
        t0 = time.perf_counter()
        ck, cd = o.ground_truth(s["data"], q[:ns], a.k, threads=cores)
        tg = time.perf_counter() - t0
        gt_cpu = {"value": ns * a.n / tg, "unit": "pairs/s", "cores": cores, "kind": "port",
                  "sample": f"{ns} of the {a.nq} queries x all {a.n} rows in {tg:.2f} s; the full batch scales linearly: {tg * a.nq / ns:.1f} s",
                  "ids_identical_to_gpu_on_sample": bool(np.array_equal(ck, s["gt_keys"][:ns]))}
        del oidx

    # ---- roofline of the search kernel: algorithmic gather bytes / kernel time --------------------------------
    X, E, Vgpu = ctr[:, 0].mean(), ctr[:, 1].mean(), ctr[:, 2].mean()
    v_note = "oracle |visited| over the whole batch"
    if v_alg is None:
        # no CPU leg (N > 1 or --no-cpu-baseline): the reference's |visited| measured on the device -- the large-capacity
        # kernel keeps the exact visited set (HB_EXACT_VISITED), so its V counter is the oracle's (asserted in the GPU tests)
        v_alg = s.get("v_exact")
        v_note = "exact visited set on the device (HB_EXACT_VISITED) over the first 1000 queries, scaled to the batch by the GPU's evaluation counts"
    V = v_alg if v_alg is not None else Vgpu
    bytes_q = 128 * V + 4 * E + 8 * X + 128 + 16 * a.k
    k_ms = float(np.mean(s["kernel_ms"]))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = bytes_q * a.nq / (k_ms * 1e-3) / 1e9
    traffic, kname = None, "hnsw_search_v2_kernel<2, true, false>"
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "search_kernel_ncu.json")))
        traffic = prof.get("dram_bytes_per_launch")
    except Exception:
        pass
    total_ms, e2e_s, e2e_pin_s = s["total_ms"], s["e2e_pageable_s"], s["e2e_pinned_s"]
    value = world * a.nq * a.steps / (total_ms * 1e-3)
    gt_ms, gt_phase = s["gt_ms"], s["gt_phase"]
    ops = 2.0 * a.nq * a.n * 1024
    tops = ops / (gt_ms * 1e-3) / 1e12
    # measured denominator: the same tcgen05.mma kind::mxf4.block_scale instruction and tile, nothing else in the kernel
    tp = C.c_double(0.0)
    if R.lib.hb_tensor_peak(R.local_rank, 0, C.byref(tp)) == 0 and tp.value > 0:
        tpeak, tpeak_src = tp.value, ("hb_tensor_peak measured in this run: M128 N240 K64 kind::mxf4.block_scale issue loop on every SM, "
                                      "no operand expansion / TMA / epilogue (of measured, per GPU x n_gpus)")
    else:
        tpeak, tpeak_src = 9000.0, "nominal dense fp4 9000 Top/s per GPU (probe failed; of nominal)"
    line = {
        "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": world, "steps": a.steps, "warmup": s["warm"],
        "ms_per_step": total_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64 xor+popcount (integer exact)", "data": "synthetic",
        "config": {"workload": workload_name(a), "recall_at_10": s["recall"], "l2": "inputs larger than L2: the index is "
                   f"{idx.info().device_bytes / 1e9:.1f} GB and one step gathers {bytes_q * a.nq / 1e9:.2f} GB",
                   "index_build_s": s["build_s"], "index_broadcast_s": s["bcast_s"] if world > 1 else None,
                   "parallelism": ("replicated index (built on rank 0, replicated GPU-to-GPU by hb_index_broadcast = ncclBroadcast inside the "
                                   "library), one query batch per GPU, no data-path collective") if world > 1 else "1 GPU",
                   "host_cores": cores,
                   "per_query": {"expansions": X, "ids_scanned": E, "visited_reference": V, "visited_reference_source": v_note, "visited_exact_on_device": s.get("v_exact"),
                                 "distance_evals_gpu": Vgpu}},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": kname, "kernel_ms": k_ms, "bytes_per_query": bytes_q,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "cpu_baseline": cpu,
        "e2e": {"value": world * a.nq * a.steps / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": a.nq * 128,
                "d2h_bytes_per_step": a.nq * a.k * 16 + a.nq * 4, "ms_per_step": e2e_s / a.steps * 1e3,
                "how": "hb_knns on PAGEABLE host buffers (a Rust Vec<Sketch> / numpy array): threaded host copy of the queries into the "
                       "index's page-locked staging, the kernel reading them and writing keys / distances / found across PCIe, threaded "
                       "copy of the results into the caller's arrays, all inside the timed region (HB_HOST_STAGING=0: cudaMemcpyAsync both ways)",
                "pinned": {"value": world * a.nq * a.steps / e2e_pin_s, "ms_per_step": e2e_pin_s / a.steps * 1e3,
                           "how": "same call on page-locked buffers: the kernel reads queries from and writes results to mapped host memory in place"}},
        "gpu_launches": int(s["launches"]), "clocks": s["clocks"], "parity": parity,
        "ground_truth_sharded": {
            "workload": f"brute-force Hamming top-{a.k}, {a.n} x {a.nq} (BASELINE.json configs[3]), rows sharded over {world} GPU(s)",
            "ms": gt_ms, "pairs_per_s": a.nq * a.n / (gt_ms * 1e-3), "checksum_sha256_16": s["gt_sha"],
            "phase_ms_rank0": {"search": gt_phase[0], "all_gather": gt_phase[1], "merge": gt_phase[2]},
            "how": "hb_ground_truth_sharded_device on every rank: tcgen05 E2M1 GEMM + fused top-k over the rank's rows, ncclAllGather of "
                   "nq*k packed (distance << 32 | key), exact merge kernel -- all inside libhnswb200; mean of 3 calls, max over ranks; "
                   "the checksum (sha256 of the packed result) is the same at every N",
            "e2e_host_buffers_ms": s["gt_e2e_ms"], "cpu_baseline": gt_cpu,
            "roofline": {"bound": "tensor", "achieved": tops, "unit": "Top/s on 4-bit operands (2*nq*N*1024 per call, whole job)",
                         "peak": tpeak * world, "frac": tops / (tpeak * world), "peak_source": tpeak_src,
                         "frac_of_nominal_9000": tops / (9000.0 * world)}},
    }
    line.update({"strong": s["extras"]} if s["extras"] else {})
    phase("report done")
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
