"""A/B of hb_knns with pageable host buffers (HB_HOST_STAGING=0..3) on the index tools/ab.py prepare parked in /dev/shm.
usage: HB_HOST_STAGING=m python tools/ab_e2e.py TAG [NQ_REP]"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import hnsw_itu_b200 as hb  # noqa: E402
from hnsw_itu_b200 import _lib  # noqa: E402

SHM = "/dev/shm/hb_ab"
meta = np.load(f"{SHM}/meta.npy")
graphs = [{"sketches": np.load(f"{SHM}/g{g}_sk.npy", mmap_mode="r"), "down": np.load(f"{SHM}/g{g}_down.npy"),
           "off": np.load(f"{SHM}/g{g}_off.npy"), "ids": np.load(f"{SHM}/g{g}_ids.npy")} for g in range(int(meta[0]))]
idx = hb.index_from_parts("hnsw", graphs, int(meta[1]), keep_host=False)
rep = int(sys.argv[2]) if len(sys.argv) > 2 else 1
q = np.ascontiguousarray(np.tile(np.load(f"{SHM}/queries.npy"), (rep, 1)))
nq, k, ef = len(q), 10, 64
keys, dists, found = np.empty((nq, k), np.uint64), np.empty((nq, k), np.uint64), np.empty(nq, np.uint32)
u64p, u32p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
args = (idx._h, q.ctypes.data_as(u64p), nq, k, ef, 0, keys.ctypes.data_as(u64p), dists.ctypes.data_as(u64p), found.ctypes.data_as(u32p), None)
lib = _lib.lib()
for _ in range(5):
    _lib.check(lib.hb_knns(*args))
ts = []
for _ in range(30):
    t0 = time.perf_counter()
    _lib.check(lib.hb_knns(*args))
    ts.append(time.perf_counter() - t0)
ts = np.array(ts) * 1e3
import hashlib  # noqa: E402
print(f"{sys.argv[1]:10s} HB_HOST_STAGING={os.environ.get('HB_HOST_STAGING', 'default')} nq={nq} median {np.median(ts):.3f} ms min {ts.min():.3f} ms "
      f"-> {nq / np.median(ts) / 1e3:.3f} MQPS  sha={hashlib.sha256(keys.tobytes() + dists.tobytes() + found.tobytes()).hexdigest()[:12]}", flush=True)
