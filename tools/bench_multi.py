"""C5 through the library-level multi-device handle (rldl_multi_*) in ONE process: refactor + solve_to_host into
pinned host arrays, wall clock around the calls.  python tools/bench_multi.py [batch] [steps]"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np, torch
import rldl_b200 as R
from rldl_b200 import mpc
import bench
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, bench.SEED), N)
mh = R.Multi(mdl, N, N, q, l, u, R.default_settings(**bench.SETTINGS))
first, dev = mh.shards()
out = {k: torch.from_numpy(v).pin_memory().numpy() for k, v in mh.empty_out().items()}
ts = []
for s in range(steps + 1):
    t0 = time.perf_counter()
    assert mh.call("refactor") == 0 and mh.solve_to_host(out) == 0
    ts.append(time.perf_counter() - t0)
it = out["iter"]
print(json.dumps({"tool": "bench_multi", "devices": sorted(set(dev)), "shards": len(first), "batch": batch,
                  "wall_ms_per_step": [round(1e3 * t, 1) for t in ts[1:]], "solves_per_s": batch / min(ts[1:]),
                  "mean_iter": float(it.mean()), "solved": float((out["status"] == 1).mean())}))
mh.free()
