"""Experiment: the latency-mode kernel with 1..6 levels of the block tree under its chain, against the streaming
kernel, over horizons that exercise every shape of leftover: python tools/exp_lat_levels.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch = 64
cfg, mdl = bench.workload("C5", batch)
bad = 0
for N in [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 16, 17, 23, 24, 25, 31, 32, 33, 47, 48, 49, 50]:
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED + N), N)
    gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**bench.SETTINGS))
    gb.set_tuning(lat_below=0); gb.solve(); ref = gb.get()
    line = f"N={N:3d} iters max {int(ref['iter'].max()):5d}:"
    for lv in range(1, 7):
        gb.set_tuning(lat_below=100000, lat_levels=lv); gb.solve(); o = gb.get()
        di = int((o["iter"] != ref["iter"]).sum()); ds = int((o["status"] != ref["status"]).sum())
        dx = float(np.abs(o["x"] - ref["x"]).max() / max(np.abs(ref["x"]).max(), 1e-12))
        line += f"  L{lv}: {di}/{ds} {dx:.1e}"
        bad += di + ds + (dx > 1e-6)
    print(line, flush=True)
    gb.free()
print("mismatches", bad)
