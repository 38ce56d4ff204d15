"""Experiment: wall-clock breakdown of the host-buffer (e2e) step of bench.py."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np, torch
import rldl_b200 as R
from rldl_b200 import mpc
import bench

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
x0 = mpc.sample_x0(mdl, batch, bench.SEED)
q, l, u = mdl.vectors(x0, N)
pin = lambda a: torch.from_numpy(a).pin_memory()
q_h, l_h, u_h = pin(q), pin(l), pin(u)
x_h = torch.empty(q.shape, dtype=torch.float64).pin_memory(); y_h = torch.empty(l.shape, dtype=torch.float64).pin_memory()
info = {k: torch.empty(batch, dtype=(torch.int64 if k in ("iter", "status", "rho_updates") else torch.float64)).pin_memory()
        for k in ("iter", "status", "obj", "pri", "dua", "rho_updates", "rho_est")}
lib = R.lib()
gb = R.Batch(mdl, N, N, q_h.numpy(), l_h.numpy(), u_h.numpy(), R.default_settings(**bench.SETTINGS))
P = lambda t: t.data_ptr()
steps = [("update_bounds", lambda: lib.rldl_batch_update_bounds(gb.h, P(l_h), P(u_h))),
         ("update_lin_cost", lambda: lib.rldl_batch_update_lin_cost(gb.h, P(q_h))),
         ("refactor", lambda: gb.refactor()), ("solve", lambda: gb.solve()),
         ("get_solution", lambda: lib.rldl_batch_get_solution(gb.h, P(x_h), P(y_h))),
         ("get_info", lambda: lib.rldl_batch_get_info(gb.h, *[P(info[k]) for k in ("iter", "status", "obj", "pri", "dua", "rho_updates", "rho_est")]))]
for rep in range(3):
    tot = 0
    for nm, fn in steps:
        torch.cuda.synchronize(); t0 = time.perf_counter(); rc = fn(); torch.cuda.synchronize(); dt = (time.perf_counter() - t0) * 1e3
        tot += dt
        print(f"rep{rep} {nm:16s} {dt:8.1f} ms rc={rc}")
    print("total", tot)
# raw copy bandwidth
d = torch.empty_like(l_h, device="cuda")
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(l_h, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("H2D GB/s", l_h.numel() * 8 / dt / 1e9)
    torch.cuda.synchronize(); t0 = time.perf_counter(); y_h.copy_(d, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("D2H GB/s", l_h.numel() * 8 / dt / 1e9)
