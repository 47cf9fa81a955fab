"""Record the reference's own run-to-run band (tests/test_sensitivity.py at a larger size) under profiles/.
usage: python tools/sensitivity.py [N] [NQ] [SEEDS] > profiles/r02_sensitivity.json   (CPU only; the oracle does the work)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc  # noqa: E402
from tests.test_sensitivity import sensitivity  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
seeds = int(sys.argv[3]) if len(sys.argv) > 3 else 8
o = orc.get()
out = {"what": "oracle re-run with randomly permuted neighbour order (what a fresh process of the reference sees: HashSet iteration, "
               "simplegraph.rs:111-117) vs the canonical .idx file order the CUDA path reproduces bit for bit",
       "n": n, "nq": nq, "k": 10, "ef": 64, "index": "m=24 M=256 efc=96, sequential CPU builder restatement", "cases": {}}
for name, planes in (("default_generator", 1024), ("tie_heavy_64_planes", 64)):
    p = orc.SynthParams(distinct_planes=planes)
    data, q = o.synth(n, 1003, p), o.synth(nq, 1004, p)
    out["cases"][name] = sensitivity(o, data, q, 10, 64, range(seeds), efc=96, m=24, M=256)
json.dump(out, sys.stdout, indent=1)
