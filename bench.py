#!/usr/bin/env python
"""bench.py — batched MPC QP solves/s through librldl_b200 (C ABI) on 1..8 B200.

Workload (BASELINE.json configs[4], "C5"): the docs quadcopter MPC (nx=12, nu=4, N=50,
box constraints, n=800, m=1412), 65536 independent instances with random x0, split
evenly across the ranks (strong scaling; instances share nothing, so there is no
collective on the solve path — torch.distributed is only the barrier / max-reduce).

One step = the hot path over the whole batch: full stage-wise KKT factorisation
(rldl_batch_refactor) + cold-start ADMM solve to convergence (rldl_batch_solve).
  value  : instances solved per second, data resident in HBM (CUDA events on the
           library's launching stream, max over ranks)
  e2e    : same through the host-buffer API: rldl_batch_update_data_refactor(q, l, u) from pinned host
           memory (= update_bounds + update_lin_cost + refactor, upload pipelined against the device work),
           rldl_batch_solve_to_host(x, y, info) into pinned host memory
--impl reference times the UNMODIFIED reference (oracle/_ref: osqp_setup_recursive +
osqp_solve_recursive) on the host cores over a bounded sample of the same instances.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))

import numpy as np  # noqa: E402

SETTINGS = dict(scaling=0, polish=0, verbose=0, time_limit=0.0, check_termination=25, eps_abs=1e-5,
                eps_rel=1e-5, eps_prim_inf=1e-4, eps_dual_inf=1e-4, alpha=1.6, sigma=1e-6, rho=0.1,
                max_iter=4000, warm_start=0, adaptive_rho=0, adaptive_rho_interval=25)
METRIC = "batched MPC QP solves/sec"
UNIT = "solves/s"
SEED = 1234


def workload(name, total_batch):
    from rldl_b200 import mpc
    cfg = dict(mpc.CONFIGS[name])
    if total_batch:
        cfg["batch"] = total_batch
    mdl = mpc.make_model(cfg)
    return cfg, mdl


def config_dict(name, cfg, n_gpus, extra=None):
    d = {"workload": f"{name}: {cfg['model']} MPC nx={cfg['nx']} nu={cfg['nu']} N={cfg['N']}, "
                     f"batch {cfg['batch']} split over {n_gpus} GPU(s), cold start, eps 1e-5, rho 0.1, "
                     f"step = full stage factorisation + ADMM solve to convergence per instance",
         "batch": cfg["batch"], "nx": cfg["nx"], "nu": cfg["nu"], "N": cfg["N"],
         "l2": "per-GPU working set (factors + iterates) is far larger than the 126 MB L2; no flush needed"}
    if extra:
        d.update(extra)
    return d


# ----------------------------------------------------------------------------
# reference arm / cpu_baseline: the unmodified reference on the host cores
# ----------------------------------------------------------------------------
def ref_sample_size(cores):
    # about 20-30 core-seconds of reference work per step (C5: ~11 ms per instance and core)
    return int(min(4096, max(256, 128 * cores)))


def parity_sample_size(cores):
    # cpu_baseline leg of the B200 arm: the reference's results on this sample are also compared with the
    # GPU results of the same instances (>= 4096 compared; ~45-90 core-seconds of reference work)
    return 8192 if cores >= 16 else 4096


def status_histogram(status):
    v, c = np.unique(np.asarray(status), return_counts=True)
    return {str(int(a)): int(b) for a, b in zip(v, c)}


def run_reference_sample(mdl, cfg, sample, seed_offset=0):
    """setup + solve of `sample` instances of the workload on all host cores (child process).
    Every completed solve counts, whatever its status (the workload holds an instance that ends at
    max_iter in the reference and on the GPU alike).
    Returns (solves_per_s, seconds, threads, outputs dict of the reference, path used)."""
    from rldl_b200 import mpc
    from oracle import ref_py
    N = cfg["N"]
    x0 = mpc.sample_x0(mdl, cfg["batch"], SEED)[seed_offset:seed_offset + sample]
    # Nmax = N+1: with N == Nmax the reference writes one element past L->p and reads the terminal
    # rho from the wrong rows; results are identical, the heap stays intact
    q, l, u = mdl.vectors(x0, N, N + 1)
    spec = dict(model=mdl, N=N, Nmax=N + 1, q=q, l=l, u=u, settings=SETTINGS, path="recursive",
                ops=[("solve",), ("get",)])
    try:
        res = ref_py.run_isolated(spec)
    except RuntimeError:
        # the reference's recursive path died in its child process (heap corruption):
        # its stock path on the assembled P/A is the same mathematics (F0) — said in the line
        spec["path"] = "stock"
        res = ref_py.run_isolated(spec)
    t = res[-1]
    secs = t["setup_wall"] + t["solve_wall"][0]
    return sample / secs, secs, t["nthreads"], res[1], spec["path"]


def parity_counts(ref, gpu, n, m):
    """Reference vs GPU on the same instances: status / iteration mismatches and the worst relative
    deviations of x, y, obj (||d||inf / max(||ref||inf, 1e-12) per instance, as the tests do) over the
    instances whose status and iteration count agree and that carry a solution."""
    rs, gs = np.asarray(ref["status"]), np.asarray(gpu["status"])
    ri, gi = np.asarray(ref["iter"]), np.asarray(gpu["iter"])
    same = (rs == gs) & (ri == gi)
    has_sol = np.isin(rs, (1, 2, -2))
    cmp_ = same & has_sol
    def rel(a, b):
        if not cmp_.any():
            return 0.0
        d = np.abs(a[cmp_] - b[cmp_]).max(axis=1)
        return float((d / np.maximum(np.abs(a[cmp_]).max(axis=1), 1e-12)).max())
    ro, go = np.asarray(ref["obj"])[cmp_], np.asarray(gpu["obj"])[cmp_]
    return {"compared": int(rs.size), "status_mismatch": int((rs != gs).sum()),
            "iter_mismatch": int(((rs == gs) & (ri != gi)).sum()),
            "max_rel_x": rel(np.asarray(ref["x"])[:, :n], np.asarray(gpu["x"])[:, :n]),
            "max_rel_y": rel(np.asarray(ref["y"])[:, :m], np.asarray(gpu["y"])[:, :m]),
            "max_rel_obj": float((np.abs(ro - go) / np.maximum(np.abs(ro), 1e-12)).max()) if ro.size else 0.0,
            "tolerance": 1e-6, "reference_status_histogram": status_histogram(rs),
            "gpu_status_histogram": status_histogram(gs)}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cfg, mdl = workload(args.workload, args.batch)
    cores = os.cpu_count() or 1
    sample = min(ref_sample_size(cores), cfg["batch"])
    for _ in range(args.warmup):
        run_reference_sample(mdl, cfg, min(sample, 4 * cores))
    tot_s, tot_n, thr, its = 0.0, 0, cores, 0.0
    hist, paths = {}, set()
    for k in range(args.steps):
        off = (k * sample) % max(1, cfg["batch"] - sample + 1)
        v, secs, thr, out, path = run_reference_sample(mdl, cfg, sample, off)
        tot_s += secs; tot_n += sample; its += float(out["iter"].mean())
        paths.add(path)
        for kk, vv in status_histogram(out["status"]).items():
            hist[kk] = hist.get(kk, 0) + vv
    value = tot_n / tot_s
    entry = "osqp_setup_recursive + osqp_solve_recursive" if paths == {"recursive"} else \
        "osqp_setup + osqp_solve on the assembled P/A where the recursive path aborted (SURVEY F4)"
    desc = (f"{sample} instances of the workload per step ({entry} each), "
            f"{thr} pthreads over instances, mean {its / max(1, args.steps):.0f} ADMM iterations, every completed "
            f"solve counted; status histogram {hist}")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(args.workload, cfg, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": thr, "kind": "reference", "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------
# the other BASELINE.json configs (parity-test cases), measured once per bench run at N = 1 outside the
# timed region: GPU throughput through the C ABI, the unmodified reference on a sample, and mismatch counts
# ----------------------------------------------------------------------------
def _ref(spec):
    from oracle import ref_py
    return ref_py.run_isolated(spec)


def scenario_c1_scalar():
    """configs[0]: one MPC QP (nx=4, nu=2, N=20) through the scalar entry points osqp_setup_recursive +
    osqp_solve_recursive — batch-of-one latency on the GPU next to the reference on one host core."""
    import ctypes as C
    import rldl_b200 as R
    from rldl_b200 import mpc
    cfg = mpc.CONFIGS["C1"]
    mdl, N = mpc.make_model(cfg), cfg["N"]
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, 1, SEED), N, N + 1)
    L, DP = R.lib(), C.POINTER(C.c_double)
    t_setup, t_solve, info = [], [], None
    for _ in range(5):
        sm = R.StageModel(mdl)
        d = sm.data
        qq, ll, uu = (np.ascontiguousarray(v[0]) for v in (q, l, u))
        d.data.n, d.data.m = len(qq), len(ll)
        d.data.q, d.data.l, d.data.u = (v.ctypes.data_as(DP) for v in (qq, ll, uu))
        w = C.POINTER(R.OSQPWorkspace)()
        st = R.default_settings(**SETTINGS)
        t0 = time.perf_counter()
        assert L.osqp_setup_recursive(C.byref(w), C.byref(d), C.byref(st), N + 1, N, mdl.nx, mdl.nu, mdl.ny, mdl.nt) == 0
        t1 = time.perf_counter()
        assert L.osqp_solve_recursive(w) == 0
        t2 = time.perf_counter()
        info = w.contents.info.contents
        it, stv, obj = int(info.iter), int(info.status_val), float(info.obj_val)
        L.osqp_cleanup(w)
        t_setup.append(t1 - t0); t_solve.append(t2 - t1)
    res = _ref(dict(model=mdl, N=N, Nmax=N + 1, q=q, l=l, u=u, settings=SETTINGS, path="recursive", nthreads=1,
                    ops=[("solve",), ("get",)]))
    r = res[1]
    return {"config": "C1: nx=4 nu=2 N=20, one instance, scalar API (osqp_setup_recursive + osqp_solve_recursive)",
            "gpu_setup_ms": 1e3 * min(t_setup), "gpu_solve_ms": 1e3 * min(t_solve),
            "reference_setup_ms": 1e3 * float(r["setup_time"][0]), "reference_solve_ms": 1e3 * float(r["solve_time"][0]),
            "iterations": it, "parity": {"compared": 1, "status_mismatch": int(stv != r["status"][0]),
                                         "iter_mismatch": int(it != r["iter"][0]),
                                         "max_rel_obj": abs(obj - r["obj"][0]) / max(abs(r["obj"][0]), 1e-12)},
            "note": "a batch of one is launch- and latency-bound on a GPU (one 384-thread CTA); the path is built for batches"}


def scenario_c2():
    import rldl_b200 as R
    from rldl_b200 import mpc
    cfg = mpc.CONFIGS["C2"]
    mdl, N, batch = mpc.make_model(cfg), cfg["N"], cfg["batch"]
    x0 = mpc.sample_x0(mdl, batch, SEED)
    q, l, u = mdl.vectors(x0, N)
    gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**SETTINGS))
    ts = []
    for _ in range(5):
        gb.refactor(); gb.solve(); ts.append(sum(gb.last_times()))
    out = gb.get(); n, m, _, _ = gb.dims(); gb.free()
    q1, l1, u1 = mdl.vectors(x0, N, N + 1)
    res = _ref(dict(model=mdl, N=N, Nmax=N + 1, q=q1, l=l1, u=u1, settings=SETTINGS, path="recursive",
                    ops=[("solve",), ("get",)]))
    tt = res[-1]
    return {"config": "C2: nx=4 nu=2 N=20, batch 1024, cold (factor + solve), device-resident",
            "gpu_ms": min(ts), "gpu_solves_per_s": batch / (min(ts) * 1e-3), "mean_iter": float(out["iter"].mean()),
            "reference_solves_per_s": batch / (tt["setup_wall"] + tt["solve_wall"][0]), "reference_threads": tt["nthreads"],
            "parity": parity_counts(res[1], out, n, m)}


def scenario_c3(sample=64):
    import rldl_b200 as R
    from rldl_b200 import mpc
    cfg = mpc.CONFIGS["C3"]
    mdl, N0, batch, T = mpc.make_model(cfg), cfg["N"], cfg["batch"], 20
    nx, nu, ny = mdl.nx, mdl.nu, mdl.ny
    x0 = mpc.sample_x0(mdl, batch, 4321)
    q, l, u = mdl.vectors(x0, N0, N0 + 1)
    idx = np.sort(np.random.default_rng(11).choice(batch, sample, replace=False))
    st = dict(SETTINGS); st["warm_start"] = 1
    import ctypes as C
    lib = R.lib()

    def closed_loop(lean):
        """lean: only x crosses PCIe per step (the feedback), as a controller would run it — the timed form;
        otherwise x, y and info of every instance are fetched for the parity comparison."""
        gb = R.Batch(mdl, N0, N0 + 1, q, l, u, R.default_settings(**st))
        xbuf = np.zeros((batch, q.shape[1]))
        N, xcur, steps, fb, k_ms = N0, x0.copy(), [], [], 0.0
        t0 = time.perf_counter()
        for t in range(T):
            assert gb.solve() == 0
            k_ms += gb.last_times()[1]
            if lean:
                if t + 1 == T:
                    break
                assert lib.rldl_batch_get_solution(gb.h, xbuf.ctypes.data_as(C.c_void_p), None) == 0
                u0 = xbuf[:, :nu]
            else:
                r = gb.get()
                n, m, _, _ = gb.dims()
                steps.append(dict(n=n, m=m, status=r["status"][idx].copy(), iter=r["iter"][idx].copy(),
                                  obj=r["obj"][idx].copy(), x=r["x"][idx, :n].copy(), y=r["y"][idx, :m].copy()))
                if t + 1 == T:
                    break
                u0 = r["x"][:, :nu]
                fb.append(u0[idx].copy())
            xcur = xcur @ mdl.A.T + u0 @ mdl.B.T
            assert gb.warm_start_shift() == 0
            Nnew = N - 1 if (t + 1) % 5 == 0 else N
            if t + 2 == T:
                Nnew = N0
            if Nnew != N:
                assert gb.update_horizon(Nnew) == 0
                N = Nnew
            dyn0 = np.ascontiguousarray(-(xcur @ mdl.A.T))
            assert gb.partial_update_bounds(ny, ny + nx, dyn0, dyn0) == 0
        wall = time.perf_counter() - t0
        gb.free()
        return wall, k_ms, steps, fb

    closed_loop(True)                        # warm-up (first-use costs of the process)
    wall, k_ms, _, _ = closed_loop(True)     # timed
    _, _, steps, fb = closed_loop(False)     # the same run with everything fetched, for the parity counts
    t0 = time.perf_counter()
    ref = _ref(dict(model=mdl, N=N0, Nmax=N0 + 1, q=q[idx], l=l[idx], u=u[idx], settings=st, path="recursive",
                    ops=[("receding", x0[idx], T, 5, fb)]))[0]
    ref_wall = time.perf_counter() - t0
    tot = {"compared": 0, "status_mismatch": 0, "iter_mismatch": 0, "max_rel_x": 0.0, "max_rel_y": 0.0, "max_rel_obj": 0.0}
    for s, rr in zip(steps, ref):
        pc = parity_counts(rr, s, s["n"], s["m"])
        for k in ("compared", "status_mismatch", "iter_mismatch"):
            tot[k] += pc[k]
        for k in ("max_rel_x", "max_rel_y", "max_rel_obj"):
            tot[k] = max(tot[k], pc[k])
    return {"config": f"C3: quadcopter nx=12 nu=4 N=50, batch {batch}, receding horizon T={T}: device-shifted warm start, "
                      f"stage-0 bound update, horizon shrink every 5th step + grow back",
            "gpu_wall_s": wall, "gpu_solves_per_s": batch * T / wall, "admm_kernel_ms_total": k_ms,
            "note": "GPU wall clock includes the host closed loop (x D2H every step, numpy plant update, stage-0 bound H2D); parity from a second run of the same loop with x, y, info of every instance fetched",
            "reference_solves_per_s": sample * T / ref_wall,
            "reference_note": f"{sample} sampled instances replayed with the GPU's feedback; wall clock includes process start",
            "parity": tot}


def scenario_c4(sample=128):
    import rldl_b200 as R
    from rldl_b200 import mpc
    cfg = mpc.CONFIGS["C4"]
    mdl, N, batch = mpc.make_model(cfg), cfg["N"], cfg["batch"]
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, SEED), N)
    rho = np.where(np.arange(batch) % 2 == 0, 1e-3, 10.0)
    st = dict(SETTINGS); st.update(adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5)
    ts = []
    for _ in range(3):
        gb = R.Batch(mdl, N, N, q, l, u, R.default_settings(**st), rho=rho)
        t0 = time.perf_counter(); assert gb.solve() == 0; ts.append(time.perf_counter() - t0)
        out = gb.get(); n, m, _, _ = gb.dims(); gb.free()
    res = _ref(dict(model=mdl, N=N, Nmax=N, q=q[:sample], l=l[:sample], u=u[:sample], rho=rho[:sample], settings=st,
                    path="stock", ops=[("solve",), ("get",)]))
    tt = res[-1]
    pc = parity_counts(res[1], {k: v[:sample] for k, v in out.items()}, n, m)
    pc["rho_updates_mismatch"] = int((res[1]["rho_updates"] != out["rho_updates"][:sample]).sum())
    return {"config": f"C4: nx=8 nu=3 N=100, batch {batch}, adaptive rho (interval 25, tolerance 1.5, rho0 in {{1e-3, 10}})",
            "gpu_ms": 1e3 * min(ts), "gpu_solves_per_s": batch / min(ts), "mean_iter": float(out["iter"].mean()),
            "mean_rho_updates": float(out["rho_updates"].mean()),
            "reference_solves_per_s": sample / tt["solve_wall"][0], "reference_threads": tt["nthreads"],
            "reference_path": "stock (rho updates segfault on the reference's recursive workspace, SURVEY F1)", "parity": pc}


def run_scenarios():
    out = {}
    for name, fn in (("C1", scenario_c1_scalar), ("C2", scenario_c2), ("C3", scenario_c3), ("C4", scenario_c4)):
        try:
            out[name] = fn()
        except Exception as e:  # a scenario must never take the bench line down
            out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return out


# ----------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.split(",") for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            r = [c.strip() for c in r]
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                       samples=len(sm))
        return out


def shard_range(rank, world, total):
    """Contiguous instance range of one rank (instances share nothing: no data-path collective)."""
    return rank * total // world, (rank + 1) * total // world


# ----------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------
def main_b200(args):
    import torch
    import rldl_b200 as R
    from rldl_b200 import mpc

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: librldl_b200 has no CPU path")
    torch.cuda.set_device(local)
    lib = R.lib()
    if lib.rldl_set_device(local):
        raise SystemExit("rldl_set_device failed")
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    cfg, mdl = workload(args.workload, args.batch)
    N, total = cfg["N"], cfg["batch"]
    lo, hi = shard_range(rank, world, total)
    mine = hi - lo
    x0 = mpc.sample_x0(mdl, total, SEED)[lo:hi]
    q_np, l_np, u_np = mdl.vectors(x0, N)
    nmax, mmax = q_np.shape[1], l_np.shape[1]
    # pinned host staging (the C ABI takes plain host pointers)
    pin = lambda a: torch.from_numpy(a).pin_memory()
    q_h, l_h, u_h = pin(q_np), pin(l_np), pin(u_np)
    x_h = torch.empty((mine, nmax), dtype=torch.float64).pin_memory()
    y_h = torch.empty((mine, mmax), dtype=torch.float64).pin_memory()
    info_h = {k: torch.empty(mine, dtype=(torch.int64 if k in ("iter", "status", "rho_updates") else torch.float64)).pin_memory()
              for k in ("iter", "status", "obj", "pri", "dua", "rho_updates", "rho_est")}
    del q_np, l_np, u_np

    gb = R.Batch(mdl, N, N, q_h.numpy(), l_h.numpy(), u_h.numpy(), R.default_settings(**SETTINGS))
    stream = torch.cuda.ExternalStream(gb.stream())
    P = lambda t: t.data_ptr()

    def step_device():
        assert gb.refactor() == 0
        assert gb.solve() == 0
        return gb.last_times() + gb.stream_stats()

    def step_e2e():
        # new q, l, u for every instance + full stage recursion: the upload runs in chunks on a second stream while
        # typing / record pass and recursion of the previous chunk run (= update_bounds + update_lin_cost + refactor)
        assert lib.rldl_batch_update_data_refactor(gb.h, P(q_h), P(l_h), P(u_h)) == 0
        # fused solve-and-deliver: x, y and info of every instance land in the pinned host arrays as the instance
        # finishes (device-initiated writes over PCIe), i.e. the D2H transfer is inside the solve
        assert lib.rldl_batch_solve_to_host(gb.h, P(x_h), P(y_h), *[P(info_h[k]) for k in
                                            ("iter", "status", "obj", "pri", "dua", "rho_updates", "rho_est")]) == 0

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        parts = [fn() for _ in range(steps)]
        e1.record(stream)
        e1.synchronize()
        barrier()
        return allmax(e0.elapsed_time(e1)), parts

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    _, _, l0 = gb.counters()
    ms_dev, parts = timed(step_device, args.steps)
    _, _, l1 = gb.counters()
    clocks = sampler.stop() if sampler else None
    launches = allsum(float(l1 - l0))

    # parity-relevant facts of the timed solves
    step_e2e()  # also refreshes info_h / warms the e2e path
    iters = info_h["iter"].numpy().astype(np.float64)
    status = info_h["status"].numpy()
    solved = allsum(float((status == 1).sum()))
    iter_sum = float(iters.sum())
    for _ in range(max(0, args.warmup - 1)):
        step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps)

    value = total * args.steps / (ms_dev * 1e-3)
    e2e_value = total * args.steps / (ms_e2e * 1e-3)
    h2d = allsum(float(8 * (l_h.numel() + u_h.numel() + q_h.numel())))
    d2h = allsum(float(8 * (x_h.numel() + y_h.numel()) + sum(t.numel() * 8 for t in info_h.values())))

    # roofline of the dominant kernel (the fused ADMM kernel): algorithmic bytes per launch
    # = sum over this rank's instances of iterations x bytes/iteration, over the
    # kernel's own CUDA-event duration measured inside the library on its launching stream
    admm_bytes, factor_bytes, _ = gb.counters()
    nx_, nu_ = cfg["nx"], cfg["nu"]
    a_ = nx_ + nu_
    # multiply-adds of one stage of the recursion with block-diagonal pivots: sweep of X^xx and X^uu,
    # W = Xinv F', S = F W, sweep of S; 2 flop each
    gb_factor_flops = 2.0 * (N * ((nx_ ** 3 + nu_ ** 3) + (nx_ ** 3 + nu_ * nu_ * nx_) + a_ * nx_ * nx_ + nx_ ** 3) + nx_ ** 3 + nu_ ** 3)
    f_ms = float(np.mean([p[0] for p in parts])); s_ms = float(np.mean([p[1] for p in parts]))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # the dominant kernel alone: CUDA events around the k_admm_stream launch and the instance-iterations that kernel
    # executed itself, both from the library (rldl_batch_stream_stats); the iterations of the latency-mode tail are
    # not counted at the streaming kernel's bytes
    st_ms = float(np.mean([p[2] for p in parts])); st_iters = float(np.mean([p[3] for p in parts]))
    if st_ms > 0:
        achieved = st_iters * admm_bytes / (st_ms * 1e-3) / 1e9
    else:  # the shard was small enough for the latency-mode kernel alone
        achieved = iter_sum * admm_bytes / (s_ms * 1e-3) / 1e9
    # DRAM traffic of the same kernel: dram__bytes_read + dram__bytes_write of its committed `ncu --set full`
    # capture at uniform load (profiles/r02_stream_ncu_full_uniform_c5_batch16384.txt: 383.68 + 65.95 GB for
    # 1.6384 M instance-iterations x 235 872 algorithmic bytes = x1.163), scaled to this launch.  Only quoted for
    # the shape and record layout the capture was taken on; null otherwise.
    traffic = None
    traffic_source = "no ncu capture for this shape / record layout"
    if abs(admm_bytes - 235872.0) < 1.0:
        traffic_ratio = (383.681287e9 + 65.951344e9) / (16384 * 100 * 235872.0)
        traffic = traffic_ratio * (st_iters if st_ms > 0 else iter_sum) * admm_bytes
        traffic_source = ("ncu dram__bytes_read+write of this kernel at batch 16384, uniform 100 iterations "
                          "(x%.3f of algorithmic), scaled by this launch's instance-iterations" % traffic_ratio)
    # FP64 peak of the box: tools/micro/dfma.cu, measured on this pool's B200 (profiles/r02_dfma_peak_b200.txt)
    fp64_peak = 33988.7e9
    flops_factor = float(gb_factor_flops) * mine
    roof = {"bound": "hbm", "kernel": "k_admm_stream (fused KKT solve + ADMM step + termination, whole solve loop; its tail is finished by k_admm_cta inside the same timed region)",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650",
            "traffic": traffic, "traffic_source": traffic_source,
            "launch_ms": st_ms if st_ms > 0 else s_ms,
            "algorithmic_bytes_per_launch": (st_iters if st_ms > 0 else iter_sum) * admm_bytes,
            "instance_iterations_in_this_kernel": st_iters, "instance_iterations_of_the_solve": iter_sum,
            "solve_ms_all_admm_kernels": s_ms,
            "note": "achieved = instance-iterations executed by k_admm_stream itself x bytes per instance-iteration / its own CUDA-event duration (rldl_batch_stream_stats); the latency-mode kernels that finish the tail (factor resident in shared memory) are inside solve_ms_all_admm_kernels but not in this figure",
            "factor_kernel": {"kernel": "k_factor_stream (stage recursion, FP64 / shared-memory bound: ~7 flop/B; profiles/r02_factor_ncu_full_c5_batch16384.txt)",
                              "launch_ms": f_ms, "achieved": factor_bytes / (f_ms * 1e-3) / 1e9,
                              "frac": factor_bytes / (f_ms * 1e-3) / 1e9 / peak,
                              "fp64_gflops": flops_factor / (f_ms * 1e-3) / 1e9,
                              "fp64_peak_gflops": fp64_peak / 1e9, "fp64_peak_source": "tools/micro/dfma.cu on B200 (profiles/r02_dfma_peak_b200.txt)",
                              "fp64_frac": flops_factor / (f_ms * 1e-3) / fp64_peak}}
    mean_iter = allsum(iter_sum) / total
    gb.free()

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": config_dict(args.workload, cfg, world),
                "workload_stats": {"mean_admm_iterations": mean_iter, "solved_fraction": solved / total},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof}
        if world == 1 and not args.no_cpu_baseline:
            # outside every timed region: the unmodified reference on the host cores over the first `sample`
            # instances — timed as the CPU baseline and used as the checker of the GPU results of the same
            # instances (x_h, y_h, info_h hold the last end-to-end step)
            cores = os.cpu_count() or 1
            sample = min(parity_sample_size(cores), total)
            v, secs, thr, ref, path = run_reference_sample(mdl, cfg, sample)
            gpu = {"status": info_h["status"].numpy()[:sample], "iter": info_h["iter"].numpy()[:sample],
                   "obj": info_h["obj"].numpy()[:sample], "x": x_h.numpy()[:sample], "y": y_h.numpy()[:sample]}
            n_act, m_act = N * (cfg["nx"] + cfg["nu"]), N * (2 * cfg["nx"] + cfg["nu"]) + cfg["nx"]
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": thr, "kind": "reference",
                                    "sample": f"first {sample} instances of the workload, "
                                              f"{'osqp_setup_recursive + osqp_solve_recursive' if path == 'recursive' else 'osqp_setup + osqp_solve (stock path)'} "
                                              f"each, {thr} pthreads, {secs:.2f} s wall, "
                                              f"mean {float(ref['iter'].mean()):.0f} ADMM iterations"}
            line["parity"] = parity_counts(ref, gpu, n_act, m_act)
            if not args.no_scenarios:
                line["scenarios"] = run_scenarios()
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C5")
    ap.add_argument("--batch", type=int, default=0, help="override the total batch (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-scenarios", action="store_true", help="skip the C1-C4 scenario measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)
    return main_b200(args)


if __name__ == "__main__":
    sys.exit(main())
