"""Experiment: per-iteration time of the latency-mode kernel alone (C5, batch = one instance per SM, fixed iteration
count) for values of one tuning knob: python tools/exp_lat_iter.py <knob> v1,v2,..."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
knob, vals = (sys.argv[1], [int(v) for v in sys.argv[2].split(",")]) if len(sys.argv) > 2 else ("lat_below", [100000])
batch, iters = 148, 2000
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
s = dict(bench.SETTINGS); s.update(max_iter=iters, eps_abs=1e-12, eps_rel=1e-12)
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
gb.set_tuning(lat_below=100000)
if os.environ.get("LAT_LOCKSTEP"): gb.set_tuning(lat_lockstep=int(os.environ["LAT_LOCKSTEP"]))
if knob != "lat_levels" and os.environ.get("LAT_LEVELS"): gb.set_tuning(lat_levels=int(os.environ["LAT_LEVELS"]))
ref = None
for rep in range(2):
    for v in vals:
        gb.set_tuning(**{knob: v})
        ts = []
        for _ in range(3):
            gb.solve(); ts.append(gb.last_times()[1])
        o = gb.get()
        if ref is None:
            ref = o
        same = bool((o["iter"] == ref["iter"]).all()) and float(np.abs(o["x"] - ref["x"]).max())
        print(f"{knob}={v}: {min(ts):.2f} ms for {iters} iterations = {1e3 * min(ts) / iters:.3f} us per iteration (max |dx| vs first {same})", flush=True)
gb.free()
