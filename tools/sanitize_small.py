"""A CI-sized pass through every kernel family for compute-sanitizer (memcheck / racecheck / synccheck): device build,
both search kernels (ef <= 256 and the large-capacity one), compact layout, the three ground-truth kernels, merge /
pack / unpack.  Results are checked against the oracle so that a sanitizer-clean run is also a correct one.
usage (GPU box): compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hnsw_itu_b200 as hb  # noqa: E402
from oracle import oracle as orc  # noqa: E402

o = orc.get()
data, q = o.synth(3000, 1003), o.synth(96, 1004)
oidx = o.build(data, "hnsw", efc=32, m=8, M=32, parallel=False)
g = hb.index_from_parts("hnsw", oidx.graphs(), oidx.ep)
for k, ef in ((10, 64), (10, 16), (20, 200), (10, 300)):
    gk, gd, gf = g.knns_arrays(q, k, ef)
    ok, od, of = oidx.knns(q, k, ef)
    assert np.array_equal(gk, ok) and np.array_equal(gd, od) and np.array_equal(gf, of), (k, ef)
g.compact()
gk, gd, gf = g.knns_arrays(q, 10, 64)
ok, od, of = oidx.knns(q, 10, 64)
assert np.array_equal(gk, ok) and np.array_equal(gd, od)
rk, rd = o.ground_truth(data, q, 10)
for algo in ("mxf4", "mma", "scan"):
    tk, td = hb.ground_truth(data, q, 10, algo=algo)
    assert np.array_equal(tk, rk) and np.array_equal(td, rd), algo
tk, td = hb.ground_truth(data, q, 100, algo="mxf4")
rk, rd = o.ground_truth(data, q, 100)
assert np.array_equal(tk, rk) and np.array_equal(td, rd)
b = hb.build_hnsw(data, efc=32, m=8, M=32, serial_prefix=500, rayon_threads=2)
bk, bd, bf = b.knns_arrays(q, 10, 64)
assert orc.recall_at_k(bk, o.ground_truth(data, q, 10)[0], 10) > 0.8
m = hb.MultiIndex([0]).replicate(g)
mk, md, mf = m.knns_arrays(q, 10, 64)
assert np.array_equal(mk, ok)
sk, sd = hb.MultiIndex([0]).ground_truth(data, q, 10)
assert np.array_equal(sk, o.ground_truth(data, q, 10)[0])
print("sanitize_small: all results identical to the oracle")
