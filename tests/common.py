"""Shared helpers for the parity tests (test infrastructure)."""
import numpy as np

import rldl_b200 as R
from rldl_b200 import mpc
from oracle import ref_py

# SURVEY §8d common settings
BASE = dict(scaling=0, polish=0, verbose=0, time_limit=0.0, check_termination=25, eps_abs=1e-5,
            eps_rel=1e-5, eps_prim_inf=1e-4, eps_dual_inf=1e-4, alpha=1.6, sigma=1e-6, rho=0.1,
            max_iter=4000, warm_start=0, adaptive_rho=0, adaptive_rho_interval=25)

REL_TOL = 1e-6  # north_star: x, y, obj within 1e-6 relative in FP64


def settings_dict(**kw):
    d = dict(BASE); d.update(kw)
    return d


def gpu_settings(**kw):
    return R.default_settings(**settings_dict(**kw))


def ref_run(mdl, N, Nmax, q, l, u, ops, path="recursive", rho=None, nthreads=None, **kw):
    """Run a scripted op sequence on the UNMODIFIED reference in a child process."""
    spec = dict(model=mdl, N=N, Nmax=Nmax, q=q, l=l, u=u, settings=settings_dict(**kw), path=path,
                rho=rho, nthreads=nthreads, ops=ops)
    try:
        return ref_py.run_isolated(spec)
    except RuntimeError as e:
        # with N == Nmax the reference's recursive path writes one element past L->p
        # (src/recursive_ldl.c:1302); whether glibc aborts the child depends on the heap layout of the
        # box.  For plain solves the stock path on the assembled P/A is the same mathematics (F0).
        plain = all(op[0] in ("solve", "get") for op in ops)
        if path == "recursive" and N == Nmax and plain and "reference worker failed" in str(e):
            spec["path"] = "stock"
            return ref_py.run_isolated(spec)
        raise


def rel_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-12)


def problem(name, batch, seed=1234, Nmax=None):
    cfg = mpc.CONFIGS[name]
    mdl = mpc.make_model(cfg)
    N = cfg["N"]
    x0 = mpc.sample_x0(mdl, batch, seed)
    q, l, u = mdl.vectors(x0, N, Nmax)
    return mdl, N, x0, q, l, u


def compare(out, ref, n, m, tag=""):
    """Per-instance parity: same status and iteration count; x, y, obj within REL_TOL."""
    st_ok = out["status"] == ref["status"]
    it_ok = out["iter"] == ref["iter"]
    bad = []
    for k in range(len(st_ok)):
        if not (st_ok[k] and it_ok[k]):
            bad.append((k, int(out["status"][k]), int(ref["status"][k]), int(out["iter"][k]), int(ref["iter"][k])))
            continue
        ex = rel_err(out["x"][k, :n], ref["x"][k, :n])
        ey = rel_err(out["y"][k, :m], ref["y"][k, :m])
        eo = abs(out["obj"][k] - ref["obj"][k]) / max(abs(ref["obj"][k]), 1e-12)
        if max(ex, ey, eo) > REL_TOL:
            bad.append((k, "x", ex, "y", ey, "obj", eo))
    return bad


def scaled_problem(name, batch, seed=1234, Nmax=None):
    """A problem whose Ruiz equilibration (reference src/scaling.c) is not the identity: the BASELINE random
    models have unit KKT column norms, so their cost / input matrices are rescaled here; every instance gets its
    own linear cost so that the cost scaling c differs inside a batch."""
    cfg = mpc.CONFIGS[name]
    mdl = mpc.make_model(cfg)
    if cfg["model"] != "quadcopter":
        mdl.Q = 40.0 * mdl.Q
        mdl.R = 0.02 * mdl.R
        mdl.B = 0.2 * mdl.B
    N = cfg["N"]
    x0 = mpc.sample_x0(mdl, batch, seed)
    q, l, u = mdl.vectors(x0, N, Nmax)
    rng = np.random.default_rng(seed + 1)
    n = N * (mdl.nx + mdl.nu)
    q[:, :n] += rng.standard_normal((batch, 1)) ** 2 * 3.0 * rng.standard_normal((batch, n))
    return mdl, N, x0, q, l, u
