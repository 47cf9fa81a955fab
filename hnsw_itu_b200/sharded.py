"""Multi-GPU host logic for the query path: one process per GPU over torch.distributed.

The reference is a single process (rayon threads); its only scaling hook is the `--start/--len` sub-range of
`index` and `ground-truth` (cli/src/main.rs:596-602, 698-704, applied at main.rs:136-140).  Two partitionings follow
from it:

  * replicated index, query batch split (BASELINE configs 2-3): every GPU holds the whole index and searches a
    contiguous slice of the batch.  Queries are independent, so there is NO data-path collective; the only
    communication is the one-off broadcast of the index arrays at load time (`broadcast_index`).
  * sharded data (ground truth, config 4; sharded HNSW, config 5): GPU r holds rows [start_r, start_r + len_r) --
    exactly what `-b start_r -l len_r` selects -- searches its shard for ALL queries, and the per-shard ascending
    (distance, key) lists are exchanged with ONE all-gather of nq*k packed u64 per rank followed by the exact merge
    kernel (hb_topk_merge_device).  The order is total (src/index/mod.rs:90-97), so the merged list equals a single
    Bruteforce::search + sort over all rows (src/index/bruteforce.rs:37-47, cli/src/main.rs:738-741).

The functions that only move tensors (`shard_bounds`, `query_slice`, `all_gather_lists`, `broadcast_graphs`) work on
any backend and are covered by world_size-2 gloo tests on CPU; everything that computes calls the CUDA library.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from ._lib import check
from .index import Index, index_from_parts

NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def shard_bounds(n: int, world: int) -> list[tuple[int, int]]:
    """Contiguous (start, len) row ranges, shard r = rows [r*n/world, (r+1)*n/world): the reference's `-b start -l len`."""
    if world < 1:
        raise ValueError("world must be >= 1")
    edges = [(r * n) // world for r in range(world + 1)]
    return [(edges[r], edges[r + 1] - edges[r]) for r in range(world)]


def query_slice(nq: int, rank: int, world: int) -> tuple[int, int]:
    """(start, len) of the query batch rank `rank` searches in replicated-index mode (results keep input order,
    src/index/mod.rs:36-41: rank r's rows land at [start, start + len) of the output)."""
    return shard_bounds(nq, world)[rank]


def _world(group=None) -> tuple[int, int]:
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def all_gather_lists(local: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather of per-shard packed lists: (nq, k) int64 on every rank -> (world, nq, k), rank-major.
    NCCL over NVLink on GPUs (0.8 MB per rank for 10 k x 10: latency bound); gloo in the CPU tests."""
    rank, world = _world(group)
    if world == 1:
        return local.unsqueeze(0)
    # concatenated-output form (dim 0 = world * nq): accepted by both NCCL and gloo
    out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local.contiguous(), group=group)
    return out.view((world,) + tuple(local.shape))


def merge_lists_device(lists: torch.Tensor) -> torch.Tensor:
    """(world, nq, k) packed ascending lists on the current CUDA device -> (nq, k) exact (distance, key) merge."""
    world, nq, k = lists.shape
    out = torch.empty((nq, k), dtype=torch.int64, device=lists.device)
    s = torch.cuda.current_stream().cuda_stream
    check(_lib.lib().hb_topk_merge_device(lists.data_ptr(), world, nq, k, out.data_ptr(), s))
    return out


def unpack_device(packed: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """packed (nq, k) -> keys, dists (int64 tensors holding u64 bit patterns; missing entries are all-ones)."""
    keys = torch.empty_like(packed)
    dists = torch.empty_like(packed)
    s = torch.cuda.current_stream().cuda_stream
    check(_lib.lib().hb_unpack_device(packed.data_ptr(), packed.numel(), keys.data_ptr(), dists.data_ptr(), s))
    return keys, dists


def ground_truth_sharded(d_shard: torch.Tensor, key_offset: int, d_queries: torch.Tensor, k: int, group=None) -> torch.Tensor:
    """CLI `ground-truth` (cli/src/main.rs:715-754) over a row-sharded data set.  d_shard (len_r, 16) int64 holds this
    rank's rows, whose first row has global key `key_offset`; d_queries (nq, 16) is the full query batch on every
    rank.  Returns the packed (nq, k) global result on every rank."""
    nq = d_queries.shape[0]
    local = torch.empty((nq, k), dtype=torch.int64, device=d_queries.device)
    s = torch.cuda.current_stream().cuda_stream
    check(_lib.lib().hb_ground_truth_device(d_shard.data_ptr(), d_shard.shape[0], key_offset, d_queries.data_ptr(), nq, k,
                                            local.data_ptr(), s))
    return merge_lists_device(all_gather_lists(local, group))


def knns_sharded(index: Index, key_offset: int, d_queries: torch.Tensor, k: int, ef: int, group=None) -> torch.Tensor:
    """Sharded HNSW (BASELINE config 5): `index` was built over this rank's row range only (local keys); every rank
    searches all queries with the same ef, local keys are shifted by `key_offset`, and the per-shard lists are
    all-gathered and merged exactly.  Returns packed (nq, k) on every rank."""
    nq = d_queries.shape[0]
    dev = d_queries.device
    keys = torch.empty((nq, k), dtype=torch.int64, device=dev)
    dists = torch.empty_like(keys)
    found = torch.empty(nq, dtype=torch.int32, device=dev)
    local = torch.empty_like(keys)
    s = torch.cuda.current_stream().cuda_stream
    index.knns_device(d_queries.data_ptr(), nq, k, ef, keys.data_ptr(), dists.data_ptr(), found.data_ptr(), 0, s)
    check(_lib.lib().hb_pack_device(keys.data_ptr(), dists.data_ptr(), found.data_ptr(), nq, k, key_offset, local.data_ptr(), s))
    return merge_lists_device(all_gather_lists(local, group))


# ---- replicated-index mode ------------------------------------------------------------------------------------

def _bcast_array(a: np.ndarray | None, shape, dtype, src: int, device, group=None) -> np.ndarray:
    """Broadcast one array as a tensor (int64/int32 views of the unsigned numpy types)."""
    tdtype = {np.dtype(np.uint64): torch.int64, np.dtype(np.uint32): torch.int32}[np.dtype(dtype)]
    sdtype = {torch.int64: np.int64, torch.int32: np.int32}[tdtype]
    if a is not None:
        t = torch.from_numpy(np.ascontiguousarray(a, dtype=dtype).view(sdtype)).to(device)
    else:
        t = torch.empty(shape, dtype=tdtype, device=device)
    dist.broadcast(t, src=src, group=group)
    return t.cpu().numpy().view(dtype)


def broadcast_graphs(graphs: list[dict] | None, ep: int | None, src: int = 0, group=None, algo: str = "hnsw") -> tuple[list[dict], int | None, str]:
    """Replicate the host graph arrays of rank `src` (Index.graph(g) dicts) on every rank.  NCCL moves them
    GPU-to-GPU over NVLink (tensors are staged on the current CUDA device); gloo moves CPU tensors."""
    rank, world = _world(group)
    if world == 1:
        return graphs, ep, algo
    device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    meta = [None]
    if rank == src:
        meta = [{"ep": ep, "algo": algo, "graphs": [{"n": int(len(g["sketches"])), "ne": int(len(g["ids"]))} for g in graphs]}]
    dist.broadcast_object_list(meta, src=src, group=group)
    out = []
    for gi, gm in enumerate(meta[0]["graphs"]):
        mine = graphs[gi] if rank == src else None
        n, ne = gm["n"], gm["ne"]
        out.append({
            "sketches": _bcast_array(mine["sketches"] if mine else None, (n, 16), np.uint64, src, device, group).reshape(n, 16),
            "down": _bcast_array(mine["down"] if mine else None, (n,), np.uint64, src, device, group),
            "off": _bcast_array(mine["off"] if mine else None, (n + 1,), np.uint64, src, device, group),
            "ids": _bcast_array(np.asarray(mine["ids"]).astype(np.uint32) if mine else None, (ne,), np.uint32, src, device, group),
        })
    return out, meta[0]["ep"], meta[0]["algo"]


def broadcast_index(index: Index | None, src: int = 0, device: int | None = None, group=None) -> Index:
    """Replicated-index mode: rank `src` holds `index` (built or loaded with keep_host); every other rank receives
    the arrays and places them in its own HBM.  Returns this rank's Index (rank `src` gets its own back)."""
    rank, world = _world(group)
    if world == 1:
        return index
    graphs = ep = None
    algo = "hnsw"
    if rank == src:
        inf = index.info()
        algo = index.algo
        graphs = [index.graph(g) for g in range(inf.n_graphs)]
        ep = None if inf.entry_point < 0 else int(inf.entry_point)
    graphs, ep, algo = broadcast_graphs(graphs, ep, src, group, algo)
    if rank == src:
        return index
    return index_from_parts(algo, graphs, ep,
                            device=torch.cuda.current_device() if device is None else device, keep_host=False)
