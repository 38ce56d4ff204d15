"""Experiment: residual trajectories at the termination checks (C5), via repeated solves with growing max_iter."""
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
pts = list(range(25, 201, 25)) + list(range(250, 1001, 50)) + list(range(1100, 4001, 100))
s = dict(bench.SETTINGS)
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**s))
pri = np.zeros((len(pts), batch), np.float32); dua = np.zeros_like(pri); its = np.zeros((len(pts), batch), np.int16)
for j, mi in enumerate(pts):
    s["max_iter"] = mi
    gb.update_settings(R.default_settings(**s))
    gb.solve()
    o = gb.get()
    pri[j] = o["pri_res"]; dua[j] = o["dua_res"]; its[j] = o["iter"]
    print(mi, (o["status"] == 1).mean(), gb.last_times()[1], flush=True)
gb.free()
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "traj_c5.npz"), pts=np.array(pts), pri=pri, dua=dua, its=its)
