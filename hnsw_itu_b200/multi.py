"""Multi-GPU host mirror over the C ABI (include/hnsw_b200.h, csrc/multi.cu).

Two launch styles, both with the collectives (ncclBroadcast / ncclAllGather + exact merge) INSIDE the library:

  * ``MultiIndex`` -- one process drives N devices (``hb_multi_*``): what a Rust/C caller of the ABI uses.
  * ``Comm`` -- one rank of a communicator for hosts that run one process per GPU (``hb_comm_*``); the unique id is
    distributed by the host side, here over ``torch.distributed``.

The reference has neither (one process, rayon threads); the partitionings mirror its ``--start/--len`` row sub-range
(cli/src/main.rs:596-602, 698-704, applied at main.rs:136-140).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import HB_PAD_LIKE_CLI, BuildParamsC, check
from .index import Index, _p, _sketches, _u32p

COMM_ID_BYTES = 128


class MultiIndex:
    """``Indexes<Sketch>`` spread over several devices of one host, driven by one process."""

    def __init__(self, devices):
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        check(_lib.lib().hb_multi_create(devs, len(devices), C.byref(h)))
        self._h = h
        self.devices = list(devices)
        self._keep = None

    def __del__(self):
        try:
            if self._h:
                _lib.lib().hb_multi_free(self._h)
                self._h = None
        except Exception:
            pass

    def size(self) -> int:
        return int(_lib.lib().hb_multi_size(self._h))

    def load_idx(self, path: str):
        """read_index (cli/src/main.rs:224-234) once, then replicate over NVLink."""
        check(_lib.lib().hb_multi_load_idx(self._h, os.fsencode(path)))
        return self

    def replicate(self, index: Index):
        """Replicate an index that lives on devices[0]; `index` is kept alive by this object."""
        check(_lib.lib().hb_multi_replicate(self._h, index._h))
        self._keep = index
        return self

    def build_sharded(self, data, efc: int = 96, m: int = 24, M: int = 256, rayon_threads: int = 16, serial_prefix: int = 0):
        """Device r builds an HNSW over rows [r*n/N, (r+1)*n/N): the reference's ``index -b start -l len`` per shard."""
        d = _sketches(data)
        bp = BuildParamsC(efc, m, M, rayon_threads, serial_prefix)
        check(_lib.lib().hb_multi_build_sharded(self._h, _p(d), len(d), C.byref(bp)))
        return self

    def load_shards(self, paths, key_offsets):
        arr = (C.c_char_p * len(paths))(*[os.fsencode(p) for p in paths])
        ko = np.ascontiguousarray(key_offsets, dtype=np.uint64)
        check(_lib.lib().hb_multi_load_shards(self._h, arr, _p(ko)))
        return self

    def knns_arrays(self, queries, k: int, ef: int, pad: bool = False):
        """Index::knns (src/index/mod.rs:28-42) over all devices: keys, dists (nq,k) u64 and found (nq) u32, input order."""
        q = _sketches(queries)
        nq = len(q)
        keys = np.zeros((nq, k), dtype=np.uint64)
        dists = np.zeros((nq, k), dtype=np.uint64)
        found = np.zeros(nq, dtype=np.uint32)
        check(_lib.lib().hb_multi_knns(self._h, _p(q), nq, k, ef, HB_PAD_LIKE_CLI if pad else 0, _p(keys), _p(dists), _p(found, _u32p)))
        return keys, dists, found

    def ground_truth(self, data, queries, k: int):
        """CLI `ground-truth` (cli/src/main.rs:715-754) with the rows sharded over the devices."""
        d, q = _sketches(data), _sketches(queries)
        keys = np.zeros((len(q), k), dtype=np.uint64)
        dists = np.zeros((len(q), k), dtype=np.uint64)
        check(_lib.lib().hb_multi_ground_truth(self._h, _p(d), len(d), _p(q), len(q), k, _p(keys), _p(dists)))
        return keys, dists

    def phase_ms(self):
        """Device time of the last sharded call on devices[0]: (search, all-gather, merge) in ms."""
        ms = (C.c_double * 3)()
        check(_lib.lib().hb_multi_phase_ms(self._h, ms))
        return tuple(ms)


class Comm:
    """One rank of the library's NCCL communicator (one process per GPU)."""

    def __init__(self, handle, rank, world):
        self._h, self.rank, self.world = handle, rank, world

    @classmethod
    def from_torch(cls, device: int, group=None) -> "Comm":
        """Create the communicator over an initialised torch.distributed group: rank 0 makes the unique id, the group
        broadcasts its 128 bytes, every rank joins."""
        import torch
        import torch.distributed as dist
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        ident = (C.c_uint8 * COMM_ID_BYTES)()
        if world > 1:
            if rank == 0:
                check(_lib.lib().hb_comm_unique_id(ident))
            on_gpu = dist.get_backend(group) == "nccl"
            t = torch.tensor(list(ident), dtype=torch.uint8, device=f"cuda:{device}" if on_gpu else "cpu")
            dist.broadcast(t, src=0, group=group)
            ident = (C.c_uint8 * COMM_ID_BYTES)(*t.cpu().tolist())
        h = C.c_void_p()
        check(_lib.lib().hb_comm_init_rank(ident, world, rank, device, C.byref(h)))
        return cls(h, rank, world)

    def __del__(self):
        try:
            if self._h:
                _lib.lib().hb_comm_free(self._h)
                self._h = None
        except Exception:
            pass

    def ground_truth_sharded(self, d_shard: int, n_local: int, key_offset: int, d_queries: int, nq: int, k: int, d_packed: int, stream: int = 0):
        check(_lib.lib().hb_ground_truth_sharded_device(self._h, d_shard, n_local, key_offset, d_queries, nq, k, d_packed, stream or None))

    def knns_sharded(self, index: Index, key_offset: int, d_queries: int, nq: int, k: int, ef: int, d_packed: int, stream: int = 0):
        check(_lib.lib().hb_knns_sharded_device(self._h, index._h, key_offset, d_queries, nq, k, ef, d_packed, stream or None))

    def broadcast_index(self, index: Index | None, root: int = 0) -> Index:
        """Replicate rank `root`'s index on every rank GPU-to-GPU (ncclBroadcast of the device arrays)."""
        out = C.c_void_p()
        check(_lib.lib().hb_index_broadcast(self._h, index._h if index is not None else None, root, C.byref(out)))
        if index is not None and self.rank == root:
            return index
        return Index(out.value)

    def phase_ms(self):
        ms = (C.c_double * 3)()
        check(_lib.lib().hb_comm_phase_ms(self._h, ms))
        return tuple(ms)
