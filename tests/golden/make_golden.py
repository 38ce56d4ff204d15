"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built by
oracle/Makefile from /root/reference).  Run here (where /root/reference exists):

    python tests/golden/make_golden.py

Every fixture stores the inputs that are not re-derivable from the seeded generator
(rldl_b200.mpc) and the reference's outputs (status, iter, obj, x, y).  The recursive path
is run with Nmax = N+1 (with N == Nmax the reference writes one element
past L->p and reads the terminal rho from the wrong rows)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "osqp-recursive-ldl_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

from common import problem, ref_run, scaled_problem, settings_dict  # noqa: E402
from rldl_b200 import mpc  # noqa: E402


def save(name, **kw):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **kw)
    print("wrote", name, {k: getattr(v, "shape", v) for k, v in kw.items()})


def cold(name, batch):
    mdl, N, x0, q, l, u = problem(name, batch, Nmax=mpc.CONFIGS[name]["N"] + 1)
    res = ref_run(mdl, N, N + 1, q, l, u, [("solve",), ("get",), ("dims",)])
    r, (n, m) = res[1], res[2]
    rng = np.random.default_rng(7)
    rhs = rng.standard_normal((2, n + m))
    kk = ref_run(mdl, N, N + 1, q, l, u, [("kkt_solve", k, rhs[k]) for k in range(2)])
    save(f"cold_{name}", batch=batch, seed=1234, n=n, m=m, status=r["status"], iter=r["iter"], obj=r["obj"],
         x=r["x"][:, :n], y=r["y"][:, :m], kkt_rhs=rhs, kkt_sol=np.stack(kk[:2]))


def adaptive(name, batch):
    """C4: adaptive rho on the reference's STOCK path (the recursive path segfaults, F1)."""
    mdl, N, x0, q, l, u = problem(name, batch)
    rho = np.where(np.arange(batch) % 2 == 0, 1e-3, 10.0)
    res = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",), ("dims",)], path="stock", rho=rho,
                  adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5)
    r, (n, m) = res[1], res[2]
    save(f"adaptive_{name}", batch=batch, seed=1234, rho=rho, n=n, m=m, status=r["status"], iter=r["iter"],
         obj=r["obj"], rho_updates=r["rho_updates"], x=r["x"][:, :n], y=r["y"][:, :m])


def receding(name, batch, T, shrink_every):
    cfg = mpc.CONFIGS[name]
    mdl, N = mpc.make_model(cfg), cfg["N"]
    x0 = mpc.sample_x0(mdl, batch, 4321)
    q, l, u = mdl.vectors(x0, N, N + 1)
    steps = ref_run(mdl, N, N + 1, q, l, u, [("receding", x0, T, shrink_every, None)], warm_start=1)[0]
    kw = dict(batch=batch, seed=4321, T=T, shrink_every=shrink_every)
    for t, s in enumerate(steps):
        for key in ("N", "iter", "status", "obj", "x", "y", "u0"):
            kw[f"{key}_{t}"] = s[key]
    save(f"receding_{name}", **kw)


def infeasible():
    """Primal infeasible: stage-1 state box that the bounded input cannot reach."""
    mdl, N, x0, q, l, u = problem("C1", 4, Nmax=21)
    b = mdl.nx + mdl.ny
    l[:, b:b + mdl.nx] = 10.0; u[:, b:b + mdl.nx] = 11.0
    r = ref_run(mdl, N, N + 1, q, l, u, [("solve",), ("get",)])[1]
    save("infeasible_C1", batch=4, seed=1234, status=r["status"], iter=r["iter"], obj=r["obj"])


def updates(name, batch):
    """Vector-update entry points between solves (osqp_update_lin_cost, osqp_partial_update_bounds,
    osqp_warm_start, osqp_warm_start_x, osqp_update_lower/upper_bound through update_bounds) on the
    recursive path, and osqp_update_rho on the stock path (the recursive path segfaults there, F1)."""
    cfg = mpc.CONFIGS[name]
    mdl, N = mpc.make_model(cfg), cfg["N"]
    x0 = mpc.sample_x0(mdl, batch, 99)
    q, l, u = mdl.vectors(x0, N, N + 1)
    rng = np.random.default_rng(5)
    n, m = N * (mdl.nx + mdl.nu), l.shape[1]
    q2 = q.copy(); q2[:, :n] += 0.05 * rng.standard_normal((batch, n))
    x0b = mpc.sample_x0(mdl, batch, 100)
    _, l2, u2 = mdl.vectors(x0b, N, N + 1)
    s, e = mdl.ny, mdl.ny + mdl.nx                    # the stage-0 dynamics rows carry x0
    xw = 0.1 * rng.standard_normal((batch, q.shape[1])); yw = 0.1 * rng.standard_normal((batch, l.shape[1]))
    lt, ut = l2.copy(), u2.copy()
    tight = slice(0, mdl.ny)                          # tighten the stage-0 path rows
    lt[:, tight] *= 0.5; ut[:, tight] *= 0.5
    ops = [("solve",), ("get",),
           ("update_lin_cost", q2), ("partial_update_bounds", s, e, l2[:, s:e].copy(), u2[:, s:e].copy()),
           ("solve",), ("get",),
           ("warm_start", xw, yw), ("solve",), ("get",),
           ("update_bounds", lt, ut), ("warm_start", xw, None), ("solve",), ("get",), ("dims",)]
    res = ref_run(mdl, N, N + 1, q, l, u, ops, warm_start=1)
    n, m = res[-2]
    kw = dict(batch=batch, seed=99, n=n, m=m, q2=q2, l2=l2, u2=u2, s=s, e=e, xw=xw, yw=yw, lt=lt, ut=ut)
    for t, i in enumerate((1, 5, 8, 12)):
        for key in ("status", "iter", "obj"):
            kw[f"{key}_{t}"] = res[i][key]
        kw[f"x_{t}"], kw[f"y_{t}"] = res[i]["x"][:, :n], res[i]["y"][:, :m]
    # stock path: per-instance osqp_update_rho between two cold solves
    rho2 = np.where(np.arange(batch) % 2 == 0, 0.02, 1.5)
    rr = ref_run(mdl, N, N, q[:, :n], l[:, :m], u[:, :m], [("solve",), ("update_rho", rho2), ("solve",), ("get",)],
                 path="stock", warm_start=0)[3]
    kw.update(rho2=rho2, status_rho=rr["status"], iter_rho=rr["iter"], obj_rho=rr["obj"],
              x_rho=rr["x"][:, :n], y_rho=rr["y"][:, :m])
    save(f"updates_{name}", **kw)


def scaled(name, batch):
    """Ruiz scaling (settings.scaling = 10, the OSQP default) on the reference's STOCK path (its recursive path
    scales the global P / A but factors the unscaled stage blocks): the equilibration itself
    (work->scaling D, E, c), a cold solve, and a solve with adaptive rho from rho = 1e-3."""
    mdl, N, x0, q, l, u = scaled_problem(name, batch)
    res = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",), ("get_scaling",), ("dims",)], path="stock", scaling=10)
    r, (D, E, c), (n, m) = res[1], res[2], res[3]
    rho = np.full(batch, 1e-3)
    ra = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",)], path="stock", rho=rho, scaling=10, adaptive_rho=1,
                 adaptive_rho_interval=25, adaptive_rho_tolerance=1.5)[1]
    save(f"scaled_{name}", batch=batch, seed=1234, n=n, m=m, D=D[:, :n], E=E[:, :m], c=c, status=r["status"],
         iter=r["iter"], obj=r["obj"], x=r["x"][:, :n], y=r["y"][:, :m], rho_a=rho, status_a=ra["status"],
         iter_a=ra["iter"], obj_a=ra["obj"], rho_updates_a=ra["rho_updates"], x_a=ra["x"][:, :n], y_a=ra["y"][:, :m])


POLISH = {"C1": (16, problem, dict(polish=1, eps_abs=3e-2, eps_rel=3e-2)),
          "C3": (8, problem, dict(polish=1, eps_abs=3e-2, eps_rel=3e-2)),
          "C4": (8, scaled_problem, dict(polish=1, eps_abs=1e-2, eps_rel=1e-2, scaling=10))}


def polished(name):
    """polish = 1 (src/polish.c) on the reference's stock path with tolerances loose enough that the guessed active
    set is sometimes wrong: accepted (status_polish 1), rejected (-1) and not attempted (0) all occur."""
    batch, prob, kw = POLISH[name]
    mdl, N, x0, q, l, u = prob(name, batch)
    res = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",), ("get_polish",), ("dims",)], path="stock", **kw)
    r, sp, (n, m) = res[1], res[2], res[3]
    save(f"polish_{name}", batch=batch, seed=1234, n=n, m=m, status=r["status"], iter=r["iter"], obj=r["obj"],
         status_polish=sp, pri_res=r["pri_res"], dua_res=r["dua_res"], x=r["x"][:, :n], y=r["y"][:, :m])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "polish":
        for nm in POLISH:
            polished(nm)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "scaled":
        scaled("C1", 6); scaled("C3", 4); scaled("C4", 4)
        sys.exit(0)
    cold("C1", 16)
    cold("C3", 6)
    cold("C4", 6)
    adaptive("C4", 8)
    receding("C1", 6, 7, 2)
    receding("C3", 3, 6, 2)
    infeasible()
    updates("C1", 6)
