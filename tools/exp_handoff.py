"""Experiment: hand-off threshold x lane on a C5 shard, with the finish-time trace: python tools/exp_handoff.py <batch>"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch = int(sys.argv[1])
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**bench.SETTINGS))
gb.set_tuning(trace=1)
gb.solve()
for lane in (8, 0):
    for below in (1024, 2048, 4096):
        gb.set_tuning(lat_below=below, lat_lane=lane)
        ts = []
        for _ in range(2):
            gb.refactor(); gb.solve(); ts.append(gb.last_times()[1])
        t, eng = gb.get_trace()
        sm, si = gb.stream_stats()
        print(f"batch {batch} lane {lane} lat_below {below}: solve {min(ts):.1f} ms; streaming kernel {sm:.1f} ms, {int((eng == 1).sum())} finished there "
              f"(last at {t[eng == 1].max():.0f} ms), latency-mode {int((eng == 2).sum())} (first at {t[eng == 2].min():.0f} ms), lane stats {gb.lane_stats()}", flush=True)
gb.free()
