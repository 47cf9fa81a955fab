"""ctypes front-end of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, bench.py's cpu_baseline /
``--impl reference`` leg and ``__graft_entry__.smoke()`` -- as the checker or
the timed CPU baseline, never by the product package ``hnsw_itu_b200``.

The functions mirror the reference's surface for this path
(``Index::search/knns`` src/index/mod.rs:21-43, ``Bruteforce`` + CLI
``ground-truth`` cli/src/main.rs:715-754, ``read_index``/``write_index``
main.rs:224-248).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)


def build(march: str = "x86-64-v3", out: str | None = None, force: bool = False) -> str:
    """Compile the C++ restatement; returns the path of the shared object."""
    out = out or os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle_capi.cpp", "hnsw_oracle.hpp")]
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(s) for s in srcs):
        return out
    subprocess.run(["make", "-C", _HERE, "-B", f"ORACLE_MARCH={march}", f"OUT={out}"], check=True, capture_output=True)
    return out


def _p(a: np.ndarray, t=_u64p):
    return a.ctypes.data_as(t)


def _sk(a) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    assert a.ndim == 2 and a.shape[1] == 16, "sketches are (n, 16) uint64"
    return a


@dataclass
class SynthParams:
    """Integer-exact SimHash-of-latent-mixture generator; frozen in BASELINE.md."""
    seed_planes: int = 1001
    seed_centres: int = 1002
    latent_dim: int = 64
    n_clusters: int = 4096
    centre_scale: int = 1
    noise_scale: int = 1
    distinct_planes: int = 1024  # 64 -> tie-stress set (every distance a multiple of 16)


class Oracle:
    def __init__(self, path: str | None = None):
        path = path or build()
        L = self.lib = C.CDLL(path)
        L.orc_last_error.restype = C.c_char_p
        L.orc_max_threads.restype = C.c_int
        L.orc_hamming.restype = C.c_uint64
        L.orc_hamming.argtypes = [_u64p, _u64p]
        L.orc_index_load_idx.restype = C.c_void_p
        L.orc_index_load_idx.argtypes = [C.c_char_p]
        L.orc_index_save_idx.argtypes = [C.c_void_p, C.c_char_p]
        L.orc_index_free.argtypes = [C.c_void_p]
        L.orc_index_size.restype = C.c_uint64
        L.orc_index_size.argtypes = [C.c_void_p]
        L.orc_index_algo.argtypes = [C.c_void_p]
        L.orc_index_num_graphs.restype = C.c_uint64
        L.orc_index_num_graphs.argtypes = [C.c_void_p]
        L.orc_index_ep.restype = C.c_int64
        L.orc_index_ep.argtypes = [C.c_void_p]
        L.orc_index_graph_info.argtypes = [C.c_void_p, C.c_uint64, _u64p, _u64p]
        L.orc_index_graph_export.argtypes = [C.c_void_p, C.c_uint64, _u64p, _u64p, _u64p, _u64p]
        L.orc_index_from_parts.restype = C.c_void_p
        L.orc_index_from_parts.argtypes = [C.c_int, C.c_uint64, _u64p, C.POINTER(_u64p), C.POINTER(_u64p), C.POINTER(_u64p),
                                           C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.orc_build.restype = C.c_void_p
        L.orc_build.argtypes = [C.c_int, _u64p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.c_int]
        L.orc_knns.argtypes = [C.c_void_p, _u64p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_int, _u64p, _u64p, _u32p, _u64p]
        L.orc_knns_bounded.argtypes = [C.c_void_p, _u64p, C.c_uint64, C.c_uint64, C.c_uint64, _u64p, _u64p, _u32p, _u64p]
        L.orc_ground_truth.argtypes = [_u64p, C.c_uint64, _u64p, C.c_uint64, C.c_uint64, C.c_int, _u64p, _u64p]
        L.orc_bruteforce_api_order.argtypes = [_u64p, C.c_uint64, _u64p, C.c_uint64, _u64p, _u64p, _u64p]
        L.orc_synth.argtypes = [C.c_uint64] * 3 + [C.c_uint32] * 5 + [C.c_uint64, C.c_uint64, _u64p]
        L.orc_golden_nsw_i32.restype = C.c_int64
        L.orc_golden_nsw_i32.argtypes = [_i32p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int32, C.c_uint64, C.c_uint64, _i32p, _u64p]
        L.orc_golden_hnsw_i32.restype = C.c_int64
        L.orc_golden_hnsw_i32.argtypes = [_i32p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int32, C.c_uint64, C.c_uint64, _i32p, _u64p]
        L.orc_golden_heuristic_i32.restype = C.c_int64
        L.orc_golden_heuristic_i32.argtypes = [_i32p, C.c_uint64, C.c_int32, C.c_uint64, _i32p]
        L.orc_golden_min_k_i32.restype = C.c_int64
        L.orc_golden_min_k_i32.argtypes = [_i32p, C.c_uint64, C.c_uint64, _i32p]
        L.orc_golden_hnsw_p3.restype = C.c_int64
        L.orc_golden_hnsw_p3.argtypes = [_i32p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, _i32p, C.c_uint64, C.c_uint64, _u64p, _u64p]
        L.orc_golden_bruteforce_p3.restype = C.c_int64
        L.orc_golden_bruteforce_p3.argtypes = [_i32p, C.c_uint64, _i32p, C.c_uint64, _u64p, _u64p]

    # -- errors ---------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.orc_last_error().decode())

    def _handle(self, h):
        if not h:
            raise RuntimeError(self.lib.orc_last_error().decode())
        return OracleIndex(self, h)

    def max_threads(self) -> int:
        return self.lib.orc_max_threads()

    # -- a1: Sketch::distance -------------------------------------------
    def hamming(self, a, b) -> int:
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = np.ascontiguousarray(b, dtype=np.uint64)
        return int(self.lib.orc_hamming(_p(a), _p(b)))

    # -- data -----------------------------------------------------------
    def synth(self, n: int, seed_rows: int, params: SynthParams = SynthParams(), start: int = 0) -> np.ndarray:
        out = np.empty((n, 16), dtype=np.uint64)
        p = params
        self.lib.orc_synth(p.seed_planes, p.seed_centres, seed_rows, p.latent_dim, p.n_clusters, p.centre_scale,
                           p.noise_scale, p.distinct_planes, start, n, _p(out))
        return out

    # -- indexes ----------------------------------------------------------
    def load_idx(self, path: str) -> "OracleIndex":
        return self._handle(self.lib.orc_index_load_idx(os.fsencode(path)))

    def build(self, data, algo: str = "hnsw", efc: int = 96, m: int = 24, M: int = 256, parallel: bool = True,
              rayon_threads: int = 16, omp_threads: int = 0) -> "OracleIndex":
        data = _sk(data)
        a = {"bruteforce": 0, "nsw": 1, "hnsw": 2}[algo]
        return self._handle(self.lib.orc_build(a, _p(data), len(data), efc, m, M, int(parallel), rayon_threads, omp_threads))

    def from_parts(self, algo: str, graphs: list[dict], ep: int | None, efc=96, m=24, M=256) -> "OracleIndex":
        """graphs[0] = base, graphs[g] = layers[g-1]; each {sketches, down?, off, ids}."""
        a = {"bruteforce": 0, "nsw": 1, "hnsw": 2}[algo]
        ng = len(graphs)
        keep = []
        ns = np.array([len(g["sketches"]) for g in graphs], dtype=np.uint64)
        sk = (_u64p * ng)()
        dn = (_u64p * ng)()
        of = (_u64p * ng)()
        ids = (C.c_void_p * ng)()
        id_bytes = 8
        for i, g in enumerate(graphs):
            s = _sk(g["sketches"]); keep.append(s); sk[i] = _p(s)
            d = np.ascontiguousarray(g.get("down", np.zeros(len(s))), dtype=np.uint64); keep.append(d); dn[i] = _p(d)
            o = np.ascontiguousarray(g.get("off", np.zeros(len(s) + 1)), dtype=np.uint64); keep.append(o); of[i] = _p(o)
            x = np.ascontiguousarray(g.get("ids", np.zeros(0, dtype=np.uint64)))
            if x.dtype == np.uint32:
                id_bytes = 4
            else:
                x = x.astype(np.uint64, copy=False)
            keep.append(x); ids[i] = x.ctypes.data
        return self._handle(self.lib.orc_index_from_parts(a, ng, _p(ns), sk, dn, of, ids, id_bytes, -1 if ep is None else ep, efc, m, M))

    # -- a7: Bruteforce / ground-truth --------------------------------------
    def ground_truth(self, data, queries, k: int, threads: int = 0):
        data, queries = _sk(data), _sk(queries)
        keys = np.zeros((len(queries), k), dtype=np.uint64)
        dists = np.zeros((len(queries), k), dtype=np.uint64)
        self._check(self.lib.orc_ground_truth(_p(data), len(data), _p(queries), len(queries), k, threads, _p(keys), _p(dists)))
        return keys, dists

    def bruteforce_api_order(self, data, query, k: int):
        data = _sk(data)
        q = np.ascontiguousarray(query, dtype=np.uint64)
        keys = np.zeros(k, dtype=np.uint64)
        dists = np.zeros(k, dtype=np.uint64)
        n = C.c_uint64(0)
        self._check(self.lib.orc_bruteforce_api_order(_p(data), len(data), _p(q), k, _p(keys), _p(dists), C.byref(n)))
        return keys[: n.value], dists[: n.value]


class OracleIndex:
    def __init__(self, o: Oracle, h):
        self.o, self.h = o, C.c_void_p(h)

    def __del__(self):
        try:
            if self.h:
                self.o.lib.orc_index_free(self.h)
                self.h = None
        except Exception:
            pass

    def size(self) -> int:
        return int(self.o.lib.orc_index_size(self.h))

    @property
    def algo(self) -> str:
        return ["bruteforce", "nsw", "hnsw"][self.o.lib.orc_index_algo(self.h)]

    @property
    def ep(self):
        e = self.o.lib.orc_index_ep(self.h)
        return None if e < 0 else int(e)

    def num_graphs(self) -> int:
        return int(self.o.lib.orc_index_num_graphs(self.h))

    def graph(self, g: int) -> dict:
        n, ne = C.c_uint64(0), C.c_uint64(0)
        self.o.lib.orc_index_graph_info(self.h, g, C.byref(n), C.byref(ne))
        sk = np.empty((n.value, 16), dtype=np.uint64)
        down = np.zeros(n.value, dtype=np.uint64)
        off = np.zeros(n.value + 1, dtype=np.uint64)
        ids = np.empty(ne.value, dtype=np.uint64)
        self.o.lib.orc_index_graph_export(self.h, g, _p(sk), _p(down), _p(off), _p(ids))
        return {"sketches": sk, "down": down, "off": off, "ids": ids}

    def graphs(self) -> list[dict]:
        return [self.graph(g) for g in range(max(1, self.num_graphs()))]

    def save_idx(self, path: str):
        self.o._check(self.o.lib.orc_index_save_idx(self.h, os.fsencode(path)))

    def knns(self, queries, k: int, ef: int, pad: bool = False, threads: int = 0, counters: bool = False):
        """Index::knns (src/index/mod.rs:28-42).  Returns keys, dists, found[, counters(X,E,V,A)]."""
        q = _sk(queries)
        nq = len(q)
        keys = np.zeros((nq, k), dtype=np.uint64)
        dists = np.zeros((nq, k), dtype=np.uint64)
        found = np.zeros(nq, dtype=np.uint32)
        ctr = np.zeros((nq, 4), dtype=np.uint64) if counters else None
        self.o._check(self.o.lib.orc_knns(self.h, _p(q), nq, k, ef, int(pad), threads, _p(keys), _p(dists), _p(found, _u32p),
                                          _p(ctr) if counters else None))
        return (keys, dists, found, ctr) if counters else (keys, dists, found)

    def knns_bounded(self, queries, k: int, ef: int):
        """Scalar model of the GPU kernel's bounded-beam formulation (keys, dists, found, counters)."""
        q = _sk(queries)
        nq = len(q)
        keys = np.zeros((nq, k), dtype=np.uint64)
        dists = np.zeros((nq, k), dtype=np.uint64)
        found = np.zeros(nq, dtype=np.uint32)
        ctr = np.zeros((nq, 4), dtype=np.uint64)
        self.o._check(self.o.lib.orc_knns_bounded(self.h, _p(q), nq, k, ef, _p(keys), _p(dists), _p(found, _u32p), _p(ctr)))
        return keys, dists, found, ctr

    def search(self, query, k: int, ef: int, pad: bool = False):
        """Index::search (mod.rs:23): list of (distance, key), ascending."""
        keys, dists, found = self.knns(np.asarray(query, dtype=np.uint64).reshape(1, 16), k, ef, pad=pad, threads=1)
        n = k if pad else int(found[0])
        return [(int(dists[0, j]), int(keys[0, j])) for j in range(n)]


def recall_at_k(ids: np.ndarray, gt: np.ndarray, k: int) -> float:
    """cli/eval/eval.py:48-56: mean |set(I[i,:k]) & set(gt[i,:k])| / k."""
    tot = 0
    for a, b in zip(ids[:, :k], gt[:, :k]):
        tot += len(set(a.tolist()) & set(b.tolist()))
    return tot / (len(ids) * k)


_default = None


def get() -> Oracle:
    global _default
    if _default is None:
        _default = Oracle()
    return _default
