"""Throughput of the BASELINE.json configs other than the bench.py workload:
C2 (C1 shape, batch 1024, cold), C3 (quadcopter, batch 4096, receding horizon T=20 with shifted warm
start, stage-0 bound updates and horizon shrink/grow), C4 (adaptive-rho stress, batch 2048) — the CUDA
path through the C ABI, and the unmodified reference (oracle/_ref, all host cores) on a sample of the
same instances.  Prints one JSON line per config.  Not the bench contract: see bench.py for that."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
from oracle import ref_py, scenarios
from common import settings_dict, gpu_settings

def sync():
    import torch; torch.cuda.synchronize()

def c2():
    cfg = mpc.CONFIGS["C1"]; mdl = mpc.make_model(cfg); N = cfg["N"]; batch = 1024
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, 1234), N)
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    ts = []
    for _ in range(5):
        gb.refactor(); gb.solve(); ts.append(sum(gb.last_times()))
    it = gb.get()["iter"]; gb.free()
    res = ref_py.run_isolated(dict(model=mdl, N=N, Nmax=N + 1, q=mdl.vectors(mpc.sample_x0(mdl, batch, 1234), N, N + 1)[0],
                                   l=mdl.vectors(mpc.sample_x0(mdl, batch, 1234), N, N + 1)[1],
                                   u=mdl.vectors(mpc.sample_x0(mdl, batch, 1234), N, N + 1)[2],
                                   settings=settings_dict(), path="recursive", ops=[("solve",), ("get",)]))
    tt = res[-1]; cpu_s = tt["setup_wall"] + tt["solve_wall"][0]
    return dict(config="C2: nx=4 nu=2 N=20, batch 1024, cold (factor + solve)", gpu_ms=min(ts), gpu_solves_per_s=batch / (min(ts) * 1e-3),
                mean_iter=float(it.mean()), cpu_solves_per_s=batch / cpu_s, cpu_threads=tt["nthreads"])

def c3(batch=4096, T=20, cpu_sample=256):
    import ctypes as C
    cfg = mpc.CONFIGS["C3"]; mdl = mpc.make_model(cfg); N0 = cfg["N"]
    nx, nu, ny = mdl.nx, mdl.nu, mdl.ny
    x0 = mpc.sample_x0(mdl, batch, 4321)
    q, l, u = mdl.vectors(x0, N0, N0)
    gb = R.Batch(mdl, N0, N0, q, l, u, gpu_settings(warm_start=1))
    lib = R.lib()
    xbuf = np.zeros((batch, q.shape[1])); N = N0; xcur = x0.copy()
    iters, ok, k_solve, k_factor = [], [], 0.0, 0.0
    sync(); t0 = time.perf_counter()
    for t in range(T):
        assert gb.solve() == 0
        k_solve += gb.last_times()[1]
        if t + 1 == T:
            break
        # closed loop through the C ABI: only x comes back (u0 = its first nu entries)
        assert lib.rldl_batch_get_solution(gb.h, xbuf.ctypes.data_as(C.c_void_p), None) == 0
        xcur = xcur @ mdl.A.T + xbuf[:, :nu] @ mdl.B.T
        assert gb.warm_start_shift() == 0
        Nnew = N - 1 if (t + 1) % 5 == 0 else N
        if t + 2 == T:
            Nnew = N0
        if Nnew != N:
            assert gb.update_horizon(Nnew) == 0   # tail re-factorisation from the saved S_k
            k_factor += gb.last_times()[0]
            N = Nnew
        dyn0 = np.ascontiguousarray(-(xcur @ mdl.A.T))   # rows ny..ny+nx of c0: l = u = -A x0
        assert gb.partial_update_bounds(ny, ny + nx, dyn0, dyn0) == 0
    sync(); wall = time.perf_counter() - t0
    out = gb.get(); gb.free()
    xs = x0[:cpu_sample]
    qs, ls, us = mdl.vectors(xs, N0, N0 + 1)
    t0 = time.perf_counter()
    ref = ref_py.run_isolated(dict(model=mdl, N=N0, Nmax=N0 + 1, q=qs, l=ls, u=us, settings=settings_dict(warm_start=1),
                                   path="recursive", ops=[("receding", xs, T, 5, None)]))
    cpu_wall = time.perf_counter() - t0
    return dict(config=f"C3: quadcopter nx=12 nu=4 N=50, batch {batch}, receding horizon T={T} (device-shifted warm start, stage-0 bound update, horizon shrink every 5th step + grow back)",
                gpu_wall_s=wall, gpu_solves_per_s=batch * T / wall, admm_kernel_ms_total=k_solve, tail_refactor_ms_total=k_factor,
                last_step_mean_iter=float(out["iter"].mean()), last_step_solved=float((out["status"] == 1).mean()),
                note="GPU wall clock includes the host-side closed loop (x D2H, numpy plant update, bound H2D)",
                cpu_solves_per_s=cpu_sample * T / cpu_wall, cpu_threads=ref[-1]["nthreads"], cpu_sample=cpu_sample,
                cpu_note="reference wall clock includes its python closed loop and process start")

def c4(batch=2048, cpu_sample=256):
    cfg = mpc.CONFIGS["C4"]; mdl = mpc.make_model(cfg); N = cfg["N"]
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, 1234), N)
    rho = np.where(np.arange(batch) % 2 == 0, 1e-3, 10.0)
    kw = dict(adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5)
    ts = []
    for _ in range(3):
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(**kw), rho=rho)
        sync(); t0 = time.perf_counter(); gb.solve(); sync(); ts.append(time.perf_counter() - t0)
        out = gb.get(); gb.free()
    res = ref_py.run_isolated(dict(model=mdl, N=N, Nmax=N, q=q[:cpu_sample], l=l[:cpu_sample], u=u[:cpu_sample], rho=rho[:cpu_sample],
                                   settings=settings_dict(**kw), path="stock", ops=[("solve",), ("get",)]))
    tt = res[-1]
    return dict(config=f"C4: nx=8 nu=3 N=100, batch {batch}, adaptive rho (interval 25, tolerance 1.5, rho0 in {{1e-3,10}})",
                gpu_solve_s=min(ts), gpu_solves_per_s=batch / min(ts), mean_iter=float(out["iter"].mean()),
                mean_rho_updates=float(out["rho_updates"].mean()), solved_fraction=float((out["status"] == 1).mean()),
                cpu_solves_per_s=cpu_sample / tt["solve_wall"][0], cpu_threads=tt["nthreads"], cpu_path="stock (F1)", cpu_sample=cpu_sample)

if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "c3", "c4"]
    for w in which:
        print(json.dumps(dict(c2=c2, c3=c3, c4=c4)[w]()), flush=True)
