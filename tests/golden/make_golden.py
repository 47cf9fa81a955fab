"""Generates the committed fixtures under tests/golden/ with the CPU oracle.

The reference (Rust) cannot be built or imported in this environment, so these vectors are
outputs of the oracle -- itself pinned to the reference's own unit-test vectors by
tests/test_oracle_golden.py.  Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as orc  # noqa: E402

CASES = [(10, 64), (10, 16), (1, 1), (10, 4), (25, 40), (100, 100)]  # (k, ef)


def main():
    o = orc.get()
    # tie-heavy on purpose: 128 distinct planes -> every distance is a multiple of 8
    p = orc.SynthParams(latent_dim=12, n_clusters=24, centre_scale=1, noise_scale=1, distinct_planes=128)
    data = o.synth(600, 7001, p)
    queries = o.synth(50, 7002, p)
    idx = o.build(data, "hnsw", efc=24, m=6, M=16, parallel=False, rayon_threads=2)
    idx.save_idx(os.path.join(HERE, "tiny_hnsw.idx"))
    nsw = o.build(data[:300], "nsw", efc=16, m=4, M=8, parallel=False)
    nsw.save_idx(os.path.join(HERE, "tiny_nsw.idx"))
    bf = o.build(data[:100], "bruteforce")
    bf.save_idx(os.path.join(HERE, "tiny_bruteforce.idx"))
    queries.tofile(os.path.join(HERE, "tiny_queries.u64"))
    out = {}
    for k, ef in CASES:
        keys, dists, found, ctr = idx.knns(queries, k, ef, counters=True)
        out[f"hnsw_k{k}_ef{ef}_keys"] = keys
        out[f"hnsw_k{k}_ef{ef}_dists"] = dists
        out[f"hnsw_k{k}_ef{ef}_found"] = found
        out[f"hnsw_k{k}_ef{ef}_ctr"] = ctr
        keys, dists, found = nsw.knns(queries, k, ef)
        out[f"nsw_k{k}_ef{ef}_keys"] = keys
        out[f"nsw_k{k}_ef{ef}_dists"] = dists
        out[f"nsw_k{k}_ef{ef}_found"] = found
    gk, gd = o.ground_truth(data, queries, 10)
    out["gt_k10_keys"], out["gt_k10_dists"] = gk, gd
    np.savez_compressed(os.path.join(HERE, "tiny_expected.npz"), **out)
    print("wrote fixtures:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
