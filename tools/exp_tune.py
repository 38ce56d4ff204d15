"""Experiment: one tuning knob of the streaming kernel swept on the same handle (C5).
python tools/exp_tune.py <batch> <max_iter> <knob> v1,v2,...   e.g.  65536 4000 l2_hints 0,1,3"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch, mi, knob = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
vals = [int(v) for v in sys.argv[4].split(",")]
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
s = dict(bench.SETTINGS); s["max_iter"] = mi
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
gb.set_tuning(**{knob: vals[0]})
gb.solve()
print('first solve done', flush=True)
for rep in range(2):
    for v in vals:
        gb.set_tuning(**{knob: v})
        ts = []
        for _ in range(2):
            gb.refactor(); gb.solve(); ts.append(gb.last_times())
        it = gb.get()["iter"]
        print(f"{knob}={v} batch {batch} max_iter {mi}: factor ms {min(t[0] for t in ts):.2f} solve ms {[round(t[1], 1) for t in ts]} "
              f"inst-iter/us {it.sum() / min(t[1] for t in ts) / 1e3:.2f} lane {gb.lane_stats()}", flush=True)
gb.free()
