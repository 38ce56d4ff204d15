"""Experiment: C5 with the OSQP default scaling = 10 (runtime-shape kernels) next to scaling = 0 (streaming kernels)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
for sc in (0, 10):
    s = dict(bench.SETTINGS); s["scaling"] = sc
    gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
    ts = []
    for _ in range(2):
        gb.refactor(); gb.solve(); ts.append(gb.last_times())
    o = gb.get(); gb.free()
    f, sv = min(t[0] for t in ts), min(t[1] for t in ts)
    print(f"scaling {sc}: batch {batch} factor {f:.1f} ms solve {sv:.1f} ms -> {batch / (f + sv) * 1e3:.0f} solves/s, mean iter {o['iter'].mean():.1f}, solved {(o['status'] == 1).mean():.4f}", flush=True)
