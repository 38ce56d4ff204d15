"""Experiment: the lockstep variant of the latency-mode kernel (a warp per item) against the group-per-item variant and
the streaming kernel, over horizons and tree levels: python tools/exp_lat_lockstep.py [workload] [batch] [N1,N2,...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cfg, mdl = bench.workload(sys.argv[1] if len(sys.argv) > 1 else "C5", batch)
bad = 0
for N in ([1, 2, 3, 4, 5, 8, 9, 16, 17, 31, 49, 50] if len(sys.argv) < 4 else [int(v) for v in sys.argv[3].split(",")]):
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED + N), N)
    gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**bench.SETTINGS))
    gb.set_tuning(lat_below=0); gb.solve(); ref = gb.get()
    line = f"N={N:3d} iters max {int(ref['iter'].max()):5d}:"
    for lk in (0, 1):
        for lv in (1, 2, 3):
            gb.set_tuning(lat_below=100000, lat_levels=lv, lat_lockstep=lk); gb.solve(); o = gb.get()
            di = int((o["iter"] != ref["iter"]).sum()); ds = int((o["status"] != ref["status"]).sum())
            dx = float(np.abs(o["x"] - ref["x"]).max() / max(np.abs(ref["x"]).max(), 1e-12))
            line += f"  K{lk}L{lv}: {di}/{ds} {dx:.1e} {gb.last_times()[1]:.2f}ms"
            bad += di + ds + (dx > 1e-6)
    print(line, flush=True)
    gb.free()
print("mismatches", bad)
