"""A/B of two builds of librldl_b200.so on the same box: python tools/exp_ab.py <lib path> <batch> <max_iter>"""
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
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
s = dict(bench.SETTINGS); s["max_iter"] = mi
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
ts = []
for _ in range(4):
    gb.solve(); ts.append(gb.last_times()[1])
it = gb.get()["iter"]
print(sys.argv[1], "batch", batch, "max_iter", mi, "solve ms", [round(t, 1) for t in ts], "inst-iter/us", it.sum() / min(ts) / 1e3)
