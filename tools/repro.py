"""Scratch reproducer: parity of the search kernel vs the oracle at a given size."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hnsw_itu_b200 as hb
from oracle import oracle as orc
N = int(sys.argv[1]); nq = int(sys.argv[2]); par = int(sys.argv[3]); efs = [int(x) for x in sys.argv[4].split(",")]
o = orc.get()
data = o.synth(N, 1003); q = o.synth(nq, 1004)
oidx = o.build(data, "hnsw", 96, 24, 256, bool(par), 16)
g = hb.index_from_parts("hnsw", oidx.graphs(), oidx.ep, keep_host=False)
print("stride", g.info().row_stride, "graphs", g.info().n_graphs, flush=True)
for ef in efs:
    gk, gd, gf, gc = g.knns_arrays(q, 10, ef, counters=True)
    ok, od, of, oc = oidx.knns(q, 10, ef, counters=True)
    print(N, nq, par, ef, "parity", np.array_equal(gk, ok) and np.array_equal(gd, od) and np.array_equal(gc.astype(np.uint64)[:,[0,1,3]], oc[:,[0,1,3]]), "V gpu/oracle", gc[:,2].mean(), oc[:, 2].mean(), flush=True)
