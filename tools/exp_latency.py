"""Experiment: latency-mode kernel vs streaming kernel at several batch sizes / hand-off thresholds (C5)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench

cfg, mdl = bench.workload("C5", 65536)
N = cfg["N"]
x0 = mpc.sample_x0(mdl, 65536, bench.SEED)
q, l, u = mdl.vectors(x0, N)
st = R.default_settings(**bench.SETTINGS)
for batch in [int(a) for a in sys.argv[1].split(",")]:
    gb = R.Batch(mdl, N, N, q[:batch], l[:batch], u[:batch], st)
    for below in sys.argv[2].split(","):
        gb.set_tuning(lat_below=int(below))
        ts = []
        for _ in range(3):
            gb.solve(); ts.append(gb.last_times()[1])
        it = gb.get()["iter"]
        print(f"batch {batch:6d} LAT_BELOW {below:>6s}: solve ms {min(ts):9.2f}  (max iter {it.max()}, sum {it.sum()}, ms/max-iter {min(ts)/it.max():.5f})", flush=True)
    gb.free()
