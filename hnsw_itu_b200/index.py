"""Host-side mirror of the reference's operator interface for the binary-sketch
query path, over the C ABI of libhnswb200.so (see include/hnsw_b200.h).

Same names, argument meaning and error behaviour as the Rust crate:
``Index.search(query, k, ef) -> [Distance]``, ``Index.knns(queries, k, ef)``,
``Index.size()``, ``Distance`` ordered by (distance, key).  All compute runs in
the CUDA library; there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from functools import total_ordering

import numpy as np

from . import _lib
from ._lib import HB_PAD_LIKE_CLI, BuildParamsC, IndexInfo, SynthParamsC, check

_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)


def _p(a: np.ndarray, t=_u64p):
    return a.ctypes.data_as(t)


def _sketches(a) -> np.ndarray:
    if isinstance(a, Sketch):
        a = a.data
    elif isinstance(a, (list, tuple)) and a and isinstance(a[0], Sketch):
        a = np.stack([s.data for s in a])
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if a.ndim == 1:
        a = a.reshape(1, -1)
    if a.ndim != 2 or a.shape[1] != 16:
        raise ValueError("sketches must have shape (n, 16) uint64")
    return a


class Sketch:
    """cli/src/sketch.rs:7-10: ``Sketch { data: [u64; 16] }``."""

    __slots__ = ("data",)

    def __init__(self, data):
        d = np.ascontiguousarray(data, dtype=np.uint64)
        if d.shape != (16,):
            raise ValueError("a Sketch has exactly 16 u64 words")
        self.data = d

    def __repr__(self):
        return f"Sketch({[int(x) for x in self.data]})"


@total_ordering
@dataclass(frozen=True)
class Distance:
    """src/index/mod.rs:49-97: ordered by (distance, key); equality compares the key only."""
    distance: int
    key: int
    point: object = None  # &'a P in the reference; a Sketch when the index keeps a host copy

    def __eq__(self, other):
        return isinstance(other, Distance) and self.key == other.key

    def __lt__(self, other):
        return (self.distance, self.key) < (other.distance, other.key)

    def __hash__(self):
        return hash(self.key)


class Index:
    """A device-resident ``Indexes<Sketch>`` (cli/src/main.rs:441-446): HNSW, NSW or Bruteforce."""

    def __init__(self, handle: int):
        self._h = C.c_void_p(handle)
        self._base_cache = None

    def __del__(self):
        try:
            if self._h:
                _lib.lib().hb_index_free(self._h)
                self._h = None
        except Exception:
            pass

    # -- Index<P> ---------------------------------------------------------------
    def size(self) -> int:
        """Index::size (src/index/mod.rs:22)."""
        return int(_lib.lib().hb_index_size(self._h))

    def info(self) -> IndexInfo:
        inf = IndexInfo()
        check(_lib.lib().hb_index_get_info(self._h, C.byref(inf)))
        return inf

    @property
    def algo(self) -> str:
        return _lib.ALGO_NAMES[self.info().algo]

    def knns_arrays(self, queries, k: int, ef: int, pad: bool = False, counters: bool = False, exact_visited: bool = False):
        """Batched search returning arrays: keys (nq,k) u64, dists (nq,k) u64, found (nq) u32
        [, counters (nq,4) u32 = X,E,V,A].  exact_visited: HB_EXACT_VISITED (the kernel with the reference's exact
        visited set whatever ef is; V then equals the reference's |visited|)."""
        q = _sketches(queries)
        nq = len(q)
        keys = np.zeros((nq, k), dtype=np.uint64)
        dists = np.zeros((nq, k), dtype=np.uint64)
        found = np.zeros(nq, dtype=np.uint32)
        ctr = np.zeros((nq, 4), dtype=np.uint32) if counters else None
        check(_lib.lib().hb_knns(self._h, _p(q), nq, k, ef, (HB_PAD_LIKE_CLI if pad else 0) | (_lib.HB_EXACT_VISITED if exact_visited else 0),
                                 _p(keys), _p(dists), _p(found, _u32p),
                                 ctr.ctypes.data_as(C.c_void_p) if counters else None))
        return (keys, dists, found, ctr) if counters else (keys, dists, found)

    def knns(self, queries, k: int, ef: int, pad: bool = False) -> list[list[Distance]]:
        """Index::knns (src/index/mod.rs:28-42): results in input order."""
        keys, dists, found = self.knns_arrays(queries, k, ef, pad=pad)
        out = []
        for i in range(len(keys)):
            n = k if pad else int(found[i])
            out.append([Distance(int(dists[i, j]), int(keys[i, j]), self._point(int(keys[i, j]))) for j in range(n)])
        return out

    def search(self, query, k: int, ef: int, pad: bool = False) -> list[Distance]:
        """Index::search (src/index/mod.rs:23)."""
        return self.knns(_sketches(query)[:1], k, ef, pad=pad)[0]

    def knns_device(self, d_queries: int, nq: int, k: int, ef: int, d_keys: int, d_dists: int, d_found: int, d_counters: int = 0,
                    stream: int = 0, pad: bool = False):
        """Device-pointer variant (raw addresses, e.g. ``tensor.data_ptr()``) on the index's device."""
        check(_lib.lib().hb_knns_device(self._h, d_queries, nq, k, ef, HB_PAD_LIKE_CLI if pad else 0, d_keys, d_dists, d_found,
                                        d_counters or None, stream or None))

    def knns_device_noalloc(self, d_queries: int, nq: int, k: int, ef: int, d_keys: int, d_dists: int, d_found: int,
                            stream: int = 0, pad: bool = False):
        """hb_knns_device_noalloc: knns_device for ef <= 256 and nq <= HB_NOALLOC_MAX_QUERIES that allocates nothing
        (scratch in a static device array shared by all calls on the device: do not overlap calls)."""
        check(_lib.lib().hb_knns_device_noalloc(self._h, d_queries, nq, k, ef, HB_PAD_LIKE_CLI if pad else 0, d_keys, d_dists, d_found,
                                                stream or None))

    # -- host copy ---------------------------------------------------------------
    def _point(self, key: int):
        if self._base_cache is None:
            try:
                self._base_cache = self.graph(0)["sketches"]
            except _lib.HbError:
                self._base_cache = False
        if self._base_cache is False:
            return None
        return Sketch(self._base_cache[key])

    def graph(self, g: int) -> dict:
        """Host copy of graph g (0 = base, g = layers[g-1]); needs keep_host."""
        n, ne = C.c_uint64(0), C.c_uint64(0)
        check(_lib.lib().hb_index_graph_dims(self._h, g, C.byref(n), C.byref(ne)))
        sk = np.empty((n.value, 16), dtype=np.uint64)
        down = np.zeros(n.value, dtype=np.uint64)
        off = np.zeros(n.value + 1, dtype=np.uint64)
        ids = np.empty(ne.value, dtype=np.uint64)
        check(_lib.lib().hb_index_graph_export(self._h, g, _p(sk), _p(down), _p(off), _p(ids)))
        return {"sketches": sk, "down": down, "off": off, "ids": ids}

    def compact(self):
        """Line-granular adjacency rows + offsets instead of fixed-stride rows (hb_index_compact): same results, ~5x less HBM."""
        check(_lib.lib().hb_index_compact(self._h))
        return self

    def write_index(self, path: str):
        """write_index (cli/src/main.rs:236-248)."""
        check(_lib.lib().hb_index_save_idx(self._h, os.fsencode(path)))


def read_index(path: str, device: int = 0, keep_host: bool = True) -> Index:
    """read_index (cli/src/main.rs:224-234): load a reference `.idx` into HBM."""
    h = C.c_void_p()
    check(_lib.lib().hb_index_load_idx(os.fsencode(path), device, int(keep_host), C.byref(h)))
    return Index(h.value)


def index_from_parts(algo: str, graphs: list[dict], ep: int | None, device: int = 0, keep_host: bool = True) -> Index:
    """graphs[0] = base, graphs[g] = layers[g-1]; each dict has sketches (n,16) u64 and, for graph indexes,
    off (n+1) u64, ids (u32 or u64, canonical neighbour order) and down (n) u64 for g > 0."""
    a = _lib.ALGO_NAMES.index(algo)
    ng = len(graphs)
    keep = []
    ns = np.array([len(g["sketches"]) for g in graphs], dtype=np.uint64)
    sk = (_u64p * ng)()
    dn = (_u64p * ng)()
    of = (_u64p * ng)()
    ids = (C.c_void_p * ng)()
    # one id width for every graph (the C ABI takes a single id_bytes): u32 only when every graph passes u32
    all_ids = [np.asarray(g.get("ids", np.zeros(0, dtype=np.uint64))) for g in graphs]
    id_bytes = 4 if all_ids and all(x.dtype == np.uint32 for x in all_ids) else 8
    id_dtype = np.uint32 if id_bytes == 4 else np.uint64
    for i, g in enumerate(graphs):
        s = _sketches(g["sketches"]) if len(g["sketches"]) else np.zeros((0, 16), dtype=np.uint64)
        d = np.ascontiguousarray(g.get("down", np.zeros(len(s))), dtype=np.uint64)
        o = np.ascontiguousarray(g.get("off", np.zeros(len(s) + 1)), dtype=np.uint64)
        x = np.ascontiguousarray(all_ids[i], dtype=id_dtype)
        keep += [s, d, o, x]
        sk[i], dn[i], of[i], ids[i] = _p(s), _p(d), _p(o), x.ctypes.data
    h = C.c_void_p()
    check(_lib.lib().hb_index_from_parts(a, ng, _p(ns), sk, dn, of, ids, id_bytes, -1 if ep is None else ep, device, int(keep_host),
                                         C.byref(h)))
    return Index(h.value)


GT_ALGOS = {"auto": 0, "scan": 1, "mma": 2, "mxf4": 3}  # hb_gt_algo


def ground_truth(data, queries, k: int, device: int = 0, algo: str = "auto"):
    """CLI `ground-truth` (cli/src/main.rs:715-754): exact k smallest (distance, key) per query,
    ascending; returns 0-based keys (nq,k) u64 and dists (nq,k) u64.  `algo`: "auto", "scan" (XOR+POPC),
    "mxf4" (tcgen05 GEMM on E2M1 operands, what "auto" picks for large batches) or "mma" (the same on int8) --
    identical results."""
    d, q = _sketches(data) if len(data) else np.zeros((0, 16), dtype=np.uint64), _sketches(queries)
    keys = np.zeros((len(q), k), dtype=np.uint64)
    dists = np.zeros((len(q), k), dtype=np.uint64)
    check(_lib.lib().hb_ground_truth_algo(_p(d), len(d), _p(q), len(q), k, device, GT_ALGOS[algo], _p(keys), _p(dists)))
    return keys, dists


def build_hnsw(data, efc: int = 96, m: int = 24, M: int = 256, rayon_threads: int = 16, device: int = 0, keep_host: bool = True,
               serial_prefix: int = 0, nsw: bool = False) -> Index:
    """HNSWBuilder::extend_parallel (src/index/hnsw.rs:44-169) with the searches on the device; nsw=True builds the
    single-graph NSW variant instead (NSWBuilder::extend_parallel, src/index/nsw.rs:224-322)."""
    d = _sketches(data)
    bp = BuildParamsC(efc, m, M, rayon_threads, serial_prefix, int(nsw))
    h = C.c_void_p()
    check(_lib.lib().hb_build_hnsw(_p(d), len(d), C.byref(bp), device, int(keep_host), C.byref(h)))
    return Index(h.value)


def synth_sketches(d_out: int, n: int, seed_rows: int, start: int = 0, seed_planes: int = 1001, seed_centres: int = 1002,
                   latent_dim: int = 64, n_clusters: int = 4096, centre_scale: int = 1, noise_scale: int = 1,
                   distinct_planes: int = 1024, stream: int = 0):
    """Fill the device buffer at address d_out with n synthetic sketches (bench/test input)."""
    sp = SynthParamsC(seed_planes, seed_centres, seed_rows, latent_dim, n_clusters, centre_scale, noise_scale, distinct_planes)
    check(_lib.lib().hb_synth_device(C.byref(sp), start, n, d_out, stream or None))
