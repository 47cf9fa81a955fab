"""A/B timing of search-kernel build variants on one index (tools/, not the bench).

  python tools/ab.py prepare N NQ            build the index once (default library), park graphs + queries in /dev/shm
  HB_LIB_PATH=variant.so python tools/ab.py run TAG [EF]   load them, time the search kernel, print QPS + result checksum
"""
import hashlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import hnsw_itu_b200 as hb  # noqa: E402

SHM = "/dev/shm/hb_ab"


def prepare(n, nq):
    os.makedirs(SHM, exist_ok=True)
    d = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(d.data_ptr(), n, 1003)
    q = torch.empty((nq, 16), dtype=torch.int64, device="cuda")
    hb.synth_sketches(q.data_ptr(), nq, 1004)
    torch.cuda.synchronize()
    t = time.time()
    idx = hb.build_hnsw(d.cpu().numpy().view(np.uint64), 96, 24, 256, rayon_threads=16, keep_host=True)
    print(f"build {time.time() - t:.1f}s", flush=True)
    inf = idx.info()
    np.save(f"{SHM}/queries.npy", q.cpu().numpy().view(np.uint64))
    meta = [inf.n_graphs, inf.entry_point]
    for g in range(inf.n_graphs):
        gr = idx.graph(g)
        np.save(f"{SHM}/g{g}_sk.npy", gr["sketches"])
        np.save(f"{SHM}/g{g}_down.npy", gr["down"])
        np.save(f"{SHM}/g{g}_off.npy", gr["off"])
        np.save(f"{SHM}/g{g}_ids.npy", gr["ids"].astype(np.uint32))
    np.save(f"{SHM}/meta.npy", np.array(meta, dtype=np.int64))


def run(tag, ef, rep=1):
    meta = np.load(f"{SHM}/meta.npy")
    graphs = [{"sketches": np.load(f"{SHM}/g{g}_sk.npy", mmap_mode="r"), "down": np.load(f"{SHM}/g{g}_down.npy"),
               "off": np.load(f"{SHM}/g{g}_off.npy"), "ids": np.load(f"{SHM}/g{g}_ids.npy")} for g in range(int(meta[0]))]
    q = np.load(f"{SHM}/queries.npy")
    q = np.tile(q, (rep, 1)) if rep >= 1 else q[:-rep]  # rep < 0: only the first -rep queries
    nq = len(q)
    idx = hb.index_from_parts("hnsw", graphs, int(meta[1]), keep_host=False)
    dq = torch.from_numpy(q.view(np.int64)).cuda()
    dk = torch.empty((nq, 10), dtype=torch.int64, device="cuda")
    dd = torch.empty_like(dk)
    df = torch.empty(nq, dtype=torch.int32, device="cuda")
    dc = torch.empty((nq, 4), dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), dc.data_ptr(), s)
    h = hashlib.sha256(dk.cpu().numpy().tobytes() + dd.cpu().numpy().tobytes()).hexdigest()[:12]
    c = dc.cpu().numpy().astype(np.float64).mean(0)
    for _ in range(3):
        idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        idx.knns_device(dq.data_ptr(), nq, 10, ef, dk.data_ptr(), dd.data_ptr(), df.data_ptr(), 0, s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{tag:28s} nq={nq} ef={ef} {ms:.3f} ms/batch {nq / ms * 1e3 / 1e6:.3f} MQPS  XEVA={c.round(1)} A_hi={(dc.cpu().numpy()[:, 3] >> 16).mean():.1f} sha={h}", flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "prepare":
        prepare(int(sys.argv[2]), int(sys.argv[3]))
    else:
        run(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 64, int(sys.argv[4]) if len(sys.argv) > 4 else 1)
