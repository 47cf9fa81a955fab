"""The GPU kernel replaces the reference's unbounded `cands` heap by a bounded
formulation (sorted beam + expanded flags + boundary-tie stack).  This checks a
scalar model of that formulation against the restatement of nsw.rs:133-184,
including tie-heavy inputs where the boundary-tie stack is exercised."""
import numpy as np
import pytest

from oracle import oracle as orc


@pytest.mark.parametrize("distinct_planes,n", [(1024, 4000), (64, 4000), (16, 3000)])
def test_bounded_model_equals_reference_semantics(oracle, distinct_planes, n):
    p = orc.SynthParams(latent_dim=16, n_clusters=64, centre_scale=1, noise_scale=1, distinct_planes=distinct_planes)
    data = oracle.synth(n, 11, p)
    q = oracle.synth(200, 12, p)
    idx = oracle.build(data, "hnsw", efc=24, m=6, M=16, parallel=True, rayon_threads=2)
    for ef in (1, 2, 3, 7, 10, 33, 64, 100):
        for k in (1, 10, 50):
            k1, d1, f1, c1 = idx.knns(q, k, ef, counters=True)
            k2, d2, f2, c2 = idx.knns_bounded(q, k, ef)
            assert np.array_equal(f1, f2)
            assert np.array_equal(k1, k2) and np.array_equal(d1, d2), (distinct_planes, ef, k)
            assert np.array_equal(c1, c2)
    if distinct_planes < 1024:  # the boundary-tie stack must actually have been used
        import ctypes
        oracle.lib.orc_debug_spill_pops.restype = ctypes.c_uint64
        assert oracle.lib.orc_debug_spill_pops() > 0


def test_bounded_model_nsw(oracle):
    p = orc.SynthParams(latent_dim=8, n_clusters=16, distinct_planes=32)
    data = oracle.synth(1500, 21, p)
    q = oracle.synth(100, 22, p)
    idx = oracle.build(data, "nsw", efc=16, m=4, M=8, parallel=False)
    for ef in (1, 5, 40):
        k1, d1, f1, c1 = idx.knns(q, 10, ef, counters=True)
        k2, d2, f2, c2 = idx.knns_bounded(q, 10, ef)
        assert np.array_equal(k1, k2) and np.array_equal(d1, d2) and np.array_equal(f1, f2) and np.array_equal(c1, c2)
