"""Experiment: residuals after the first few termination checks vs final iteration count (C5)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
x0 = mpc.sample_x0(mdl, batch, bench.SEED)
q, l, u = mdl.vectors(x0, N)
res = {}
for mi in (25, 50, 75, 100, 150, 200, 4000):
    s = dict(bench.SETTINGS); s["max_iter"] = mi
    gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
    gb.solve()
    o = gb.get(); gb.free()
    res[f"pri{mi}"] = o["pri_res"]; res[f"dua{mi}"] = o["dua_res"]; res[f"it{mi}"] = o["iter"].astype(np.int32)
    res[f"st{mi}"] = o["status"].astype(np.int32)
    print(mi, "done", (o["status"] == 1).mean(), flush=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "predict_c5.npz"), **res)
