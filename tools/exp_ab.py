"""Uniform-load run for ncu captures / A-B of builds: python tools/exp_ab.py <lib path|default> <batch> <max_iter> [scaling]
(every instance runs exactly max_iter iterations when max_iter is below the first convergences)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
if sys.argv[1] != "default":
    R.LIB_PATH = os.path.join(ROOT, sys.argv[1])
from rldl_b200 import mpc
import bench
batch = int(sys.argv[2]); mi = int(sys.argv[3])
scaling = int(sys.argv[4]) if len(sys.argv) > 4 else 0
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
s = dict(bench.SETTINGS); s["max_iter"] = mi; s["scaling"] = scaling
if scaling:
    s["eps_abs"] = s["eps_rel"] = 1e-12   # nobody converges before max_iter: uniform load
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
gb.set_tuning(lat_below=0, lat_lane=0)
ts = []
for _ in range(4):
    gb.solve(); ts.append(gb.last_times()[1])
it = gb.get()["iter"]
print(sys.argv[1], "batch", batch, "max_iter", mi, "scaling", scaling, "solve ms", [round(t, 1) for t in ts], "inst-iter/us", it.sum() / min(ts) / 1e3,
      "bytes per instance-iteration", gb.counters()[0], "stream stats", gb.stream_stats())
