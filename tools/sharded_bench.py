"""BASELINE.json configs[3] and [4] on N GPUs of one node (tools/, not the bench):
  configs[3]  brute-force ground truth, data rows sharded N ways, per-shard tcgen05 top-k, NCCL all-gather + exact merge
  configs[4]  sharded HNSW: one index per row range (the reference's `index -b start -l len`), every rank searches all
              queries, NCCL all-gather + exact merge; recall against the sharded ground truth
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/sharded_bench.py [N_TOTAL] [NQ] [K] [EF] [gt-only]
Times are CUDA events on each rank, max over ranks; results are checked against the single-GPU call on rank 0's view
when the whole data set fits (N_TOTAL <= 20 M)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import hnsw_itu_b200 as hb
from hnsw_itu_b200 import _lib, sharded

NT = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
NQ = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 10
EF = int(sys.argv[4]) if len(sys.argv) > 4 else 64
GT_ONLY = len(sys.argv) > 5 and sys.argv[5] == "gt-only"   # configs[3] alone: skip the per-shard index build
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
L = _lib.lib()

start, ln = sharded.shard_bounds(NT, world)[rank]
d = torch.empty((ln, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(d.data_ptr(), ln, 1003, start=start)          # the same global rows whatever the sharding
q = torch.empty((NQ, 16), dtype=torch.int64, device="cuda")
hb.synth_sketches(q.data_ptr(), NQ, 1004)
torch.cuda.synchronize()


def timed(fn, reps=3):
    best = 1e30
    out = None
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    return best, out


res = {"n_total": NT, "nq": NQ, "k": K, "ef": EF, "n_gpus": world, "rows_per_rank": ln}
gt_ms, gt = timed(lambda: sharded.ground_truth_sharded(d, start, q, K), reps=5 if GT_ONLY else 3)
res["ground_truth"] = {"ms": gt_ms, "pairs_per_s": NT * NQ / gt_ms * 1e3, "Top_per_s_all_gpus": 2.0 * NT * NQ * 1024 / gt_ms * 1e3 / 1e12,
                       "collective": "all_gather_into_tensor of nq*k u64 per rank + topk_merge_kernel" if world > 1 else "none"}
gtk = gt.cpu().numpy().view(np.uint64)
res["ground_truth"]["sha"] = __import__("hashlib").sha256(gtk.tobytes()).hexdigest()[:16]   # equal across --nproc-per-node values

if GT_ONLY:
    if rank == 0:
        print(json.dumps(res), flush=True)
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0)
# sharded HNSW over this rank's rows
t0 = time.time()
idx = hb.build_hnsw(d.cpu().numpy().view(np.uint64), 96, 24, 256, rayon_threads=16, keep_host=False, device=lr)
res["index_build_s"] = time.time() - t0
ms, packed = timed(lambda: sharded.knns_sharded(idx, start, q, K, EF))
keys = packed.cpu().numpy().view(np.uint64) & np.uint64(0xFFFFFFFF)
gkeys = gtk & np.uint64(0xFFFFFFFF)
recall = float(np.mean([len(set(a.tolist()) & set(b.tolist())) for a, b in zip(keys, gkeys)]) / K)
res["sharded_hnsw"] = {"ms": ms, "qps": NQ / ms * 1e3, "recall_at_k": recall,
                       "note": "every rank searches all queries on its shard's index; QPS is batch / max-over-ranks time"}
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
