"""Experiment: who finishes which instance, and when (C5 shard): python tools/exp_trace.py <batch> <lat_lane> [sched_key [retire_below]]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch, lane = int(sys.argv[1]), int(sys.argv[2])
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**bench.SETTINGS))
gb.set_tuning(lat_lane=lane, trace=1)
if len(sys.argv) > 3:
    gb.set_tuning(sched_key=int(sys.argv[3]))
if len(sys.argv) > 4:
    gb.set_tuning(lat_retire_below=int(sys.argv[4]))
for _ in range(2):
    gb.refactor(); gb.solve()
t, eng = gb.get_trace()
it = gb.get()["iter"]
print(f"batch {batch} lane {lane} args {sys.argv[3:]}: solve {gb.last_times()[1]:.1f} ms, lane stats {gb.lane_stats()}")
print("finished by streaming kernel:", int((eng == 1).sum()), "last at %.1f ms" % t[eng == 1].max(),
      "| by latency-mode CTAs:", int((eng == 2).sum()), "first at %.1f ms, last at %.1f ms" % (t[eng == 2].min(), t[eng == 2].max()))
o = np.argsort(-it)[:24]
print("longest instances: iter / engine / finish ms")
print(" ".join(f"{it[i]}/{eng[i]}/{t[i]:.0f}" for i in o))
lat = np.sort(t[eng == 2])
print("latency-mode finish times, deciles:", np.round(np.percentile(lat, [0, 10, 25, 50, 75, 90, 100]), 1))
late = np.argsort(-t)[:16]
print("last to finish: iter / engine / ms:", " ".join(f"{it[i]}/{eng[i]}/{t[i]:.0f}" for i in late))
gb.free()
