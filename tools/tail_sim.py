"""Why the 10 k-query batch stops at ~0.45 of the copy bandwidth: list-scheduling model of the batch tail.

Per-query cost = V (distance evaluations) of the oracle's search on an oracle-built index (CPU only, no GPU needed).
Queries are handed to `slots` warp contexts in arrival order the way the kernel's atomic counter does; the makespan over
the work-conserving ideal (sum / slots) is the tail's price at constant per-context speed, an upper bound since the
machine speeds up per context as it drains.  Also: longest-first order with a perfect and with noisy cost predictors.
usage: python tools/tail_sim.py [N [NQ]] > profiles/r02_tail_model.json"""
import heapq
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
o = orc.get()
idx = o.build(o.synth(n, 1003))
cost = idx.knns(o.synth(nq, 1004), 10, 64, counters=True)[3][:, 2].astype(np.float64)


def makespan(order, slots):
    h = [0.0] * slots
    end = 0.0
    for i in order:
        t = heapq.heappop(h) + cost[i]
        end = max(end, t)
        heapq.heappush(h, t)
    return end


rng = np.random.default_rng(0)
z = (cost - cost.mean()) / cost.std()
out = {"n": n, "nq": nq, "V_mean": cost.mean(), "V_cv": cost.std() / cost.mean(), "V_p99_over_mean": np.percentile(cost, 99) / cost.mean(),
       "V_max_over_mean": cost.max() / cost.mean(), "slots": {}}
for slots in (4144, 5328, 7104, 10000):
    ideal = cost.sum() / slots
    row = {"queries_per_slot": nq / slots, "arrival_order": makespan(range(nq), slots) / ideal, "longest_first_perfect": makespan(np.argsort(-cost), slots) / ideal}
    for rho in (0.3, 0.5, 0.7, 0.9):
        pred = rho * z + np.sqrt(1 - rho * rho) * rng.standard_normal(nq)
        row[f"longest_first_rho_{rho}"] = makespan(np.argsort(-pred), slots) / ideal
    out["slots"][str(slots)] = row
print(json.dumps(out, indent=1, default=float))
