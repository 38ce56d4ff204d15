"""Runs the hot paths once and prints a digest per path (TEST INFRASTRUCTURE; used by tests/test_debug_gpu.py with
RLDL_B200_LIB pointing at the -DRLDL_DEBUG build, whose in-kernel asserts trap on any out-of-range ring slot,
record offset, scheduler slot or instance id)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "osqp-recursive-ldl_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
from common import gpu_settings, problem

print("LIB", os.path.basename(R.LIB_PATH))


def digest(tag, out, n, m):
    print(f"DIGEST {tag} status {np.bincount(out['status'] + 10).tolist()} iters {int(out['iter'].sum())} "
          f"x {np.abs(out['x'][:, :n]).sum():.9e} y {np.abs(out['y'][:, :m]).sum():.9e}", flush=True)


# streaming kernel + scheduler + lane + tail hand-off (C1 shape, batch above the resident groups)
mdl, N, x0, q, l, u = problem("C1", 12000)
gb = R.Batch(mdl, N, N + 2, *mdl.vectors(x0, N, N + 2), gpu_settings())
gb.set_tuning(lat_below=1024, lat_lane=8)
assert gb.solve() == 0
n, m, _, _ = gb.dims()
digest("C1_stream_lane_tail", gb.get(), n, m)
# horizon shrink / grow (tail re-factorisation), shifted warm start, partial bound update
assert gb.warm_start_shift() == 0 and gb.update_horizon(N - 1) == 0
assert gb.update_settings(gpu_settings(warm_start=1)) == 0 and gb.solve() == 0
n, m, _, _ = gb.dims()
digest("C1_shrink_warm", gb.get(), n, m)
assert gb.update_horizon(N + 2) == 0 and gb.solve() == 0
n, m, _, _ = gb.dims()
digest("C1_grow", gb.get(), n, m)
gb.free()
# C5 shape: hand-off inside a mid-size batch, latency-mode kernel alone, streaming alone
mdl, N, x0, q, l, u = problem("C3", 600)
for tag, below in (("handoff", 200), ("latency", 100000), ("stream", 0)):
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    gb.set_tuning(lat_below=below)
    assert gb.solve() == 0
    n, m, _, _ = gb.dims()
    digest("C3_" + tag, gb.get(), n, m)
    gb.free()
# adaptive rho: park / batched refactor of flagged instances / resume
mdl, N, x0, q, l, u = problem("C4", 96)
rho = np.where(np.arange(96) % 2 == 0, 1e-3, 10.0)
gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5), rho=rho)
gb.set_tuning(lat_below=24)
assert gb.solve() == 0
n, m, _, _ = gb.dims()
digest("C4_adaptive", gb.get(), n, m)
gb.free()
print("PATHS DONE")
