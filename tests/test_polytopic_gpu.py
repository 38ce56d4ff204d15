"""General stage blocks through the C ABI: polytopic path rows touching several variables, ny != nx+nu,
nt != nx (also nt > ny), a stage cost with state-input cross terms — everything the reference's
`osqp_setup_recursive(…, nx, nu, ny, nt)` admits (include/recursive_ldl.h:17-50, 63-65) beyond the
box-constrained BASELINE shapes.  These models are outside the specialisation table, so they run on the
runtime-shape kernels (rldl_kernels.cu).  The initial states are drawn wide enough that the batch mixes
solved instances (50 … 1 000+ iterations), primal-infeasible ones and ones that stop at max_iter.

Oracle: the unmodified reference's stock path on the assembled P/A for every shape, and its recursive
path as well where that is sound (it types the terminal rows from the first nt rows of
stage Nmax-1, so it is wrong for nt > ny — it reports "non convex" where the stock path solves in 50
iterations — and for a loose row among those).
"""
import dataclasses

import numpy as np
import pytest

import rldl_b200 as R
from rldl_b200 import mpc
from common import gpu_settings, ref_run, compare, rel_err, REL_TOL

pytestmark = pytest.mark.gpu

SHAPES = [  # nx, nu, ny, nt, N, Nmax, reference recursive path sound?
    (3, 2, 4, 2, 8, 10, True),
    (4, 2, 6, 4, 12, 12, True),   # the C1 block sizes, but C is dense: not served by the specialised kernels
    (5, 3, 7, 3, 6, 8, True),
    (5, 3, 3, 6, 6, 6, False),    # nt > ny
]


def _problem(nx, nu, ny, nt, N, Nmax, batch=16, scale=3.0, seed=7):
    mdl = dataclasses.replace(mpc.polytopic_model(nx, nu, ny, nt), x0_scale=scale)
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, seed), N, Nmax)
    return mdl, q, l, u


def _check(out, ref, n, m):
    """status and iterations of every instance; x, y, obj where the reference returns a solution"""
    assert (out["status"] == ref["status"]).all(), (out["status"], ref["status"])
    assert (out["iter"] == ref["iter"]).all(), (out["iter"], ref["iter"])
    sol = np.isin(ref["status"], (1, 2, -2))
    sub = lambda d: {k: v[sol] for k, v in d.items() if k in ("status", "iter", "x", "y", "obj")}
    bad = compare(sub(out), sub(ref), n, m)
    assert not bad, bad[:3]
    return sol


@pytest.mark.parametrize("nx,nu,ny,nt,N,Nmax,rec_ok", SHAPES)
def test_polytopic_cold_solve_matches_reference(nx, nu, ny, nt, N, Nmax, rec_ok):
    mdl, q, l, u = _problem(nx, nu, ny, nt, N, Nmax)
    qs, ls, us = mdl.vectors(mpc.sample_x0(mdl, 16, 7), N, N)
    ref = ref_run(mdl, N, N, qs, ls, us, [("solve",), ("get",)], path="stock")[1]
    gb = R.Batch(mdl, N, Nmax, q, l, u, gpu_settings())
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    assert (n, m) == (N * (nx + nu), N * (nx + ny) + nt)
    sol = _check(out, ref, n, m)
    assert sol.sum() >= 8 and len(np.unique(ref["iter"])) >= 3   # the batch is not trivially uniform
    if rec_ok:
        rec = ref_run(mdl, N, Nmax, q, l, u, [("solve",), ("get",)], path="recursive")[1]
        _check(out, rec, n, m)
    gb.free()


@pytest.mark.parametrize("nx,nu,ny,nt,N,Nmax,rec_ok", SHAPES[:3])
def test_polytopic_kkt_solve_matches_dense_kkt(nx, nu, ny, nt, N, Nmax, rec_ok):
    """`linsys_solver->solve` (qdldl_interface.c:559-585) on a general stage model against a dense solve of the
    assembled KKT matrix: [x_tilde; z_tilde] with z_tilde = rhs_z + rho^-1 nu."""
    mdl, q, l, u = _problem(nx, nu, ny, nt, N, Nmax, batch=4, scale=1.0)
    s = gpu_settings()
    gb = R.Batch(mdl, N, Nmax, q, l, u, s)
    n, m, _, _ = gb.dims()
    P, A = mdl.assemble(N)
    Pf = (P + P.T).toarray() - np.diag(P.diagonal())
    Ad = A.toarray()
    rng = np.random.default_rng(5)
    rhs = rng.standard_normal((4, n + m))
    sol = gb.kkt_solve(rhs)
    for k in range(4):
        lo, up = l[k, :m], u[k, :m]
        rho = np.where((lo < -1e26) & (up > 1e26), 1e-6, np.where(up - lo < 1e-4, 1e3 * s.rho, s.rho))
        K = np.block([[Pf + s.sigma * np.eye(n), Ad.T], [Ad, -np.diag(1.0 / rho)]])
        r = rhs[k]
        w = np.linalg.solve(K, r)
        zt = r[n:] + w[n:] / rho
        assert rel_err(sol[k, :n], w[:n]) < REL_TOL
        assert rel_err(sol[k, n:], zt) < REL_TOL
    gb.free()


def test_polytopic_horizon_change_and_adaptive_rho():
    """Tail re-factorisation (LDL_update_from_pivot, src/recursive_ldl.c:946-1110) and rho-change
    re-factorisation on a general stage model: shrink, grow back, and an adaptive-rho solve, each against a
    fresh stock-path setup of the reference at that horizon."""
    nx, nu, ny, nt, N, Nmax = 5, 3, 7, 3, 6, 8
    mdl, q, l, u = _problem(nx, nu, ny, nt, N, Nmax, batch=12, scale=2.0)
    x0 = mpc.sample_x0(mdl, 12, 7)
    gb = R.Batch(mdl, N, Nmax, q, l, u, gpu_settings())
    for Nn in (4, 8, 6):
        assert gb.update_horizon(Nn) == 0
        qn, ln, un = mdl.vectors(x0, Nn, Nmax)
        assert gb.update_bounds(ln, un) == 0      # the terminal rows move with the horizon
        assert gb.solve() == 0
        out = gb.get()
        n, m, Ncur, _ = gb.dims()
        assert Ncur == Nn
        qs, ls, us = mdl.vectors(x0, Nn, Nn)
        ref = ref_run(mdl, Nn, Nn, qs, ls, us, [("solve",), ("get",)], path="stock")[1]
        _check(out, ref, n, m)
    gb.free()
    # adaptive rho (stock path is the only reference that has it)
    kw = dict(adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5)
    rho0 = np.where(np.arange(12) % 2 == 0, 1e-3, 10.0)
    qs, ls, us = mdl.vectors(x0, N, N)
    ref = ref_run(mdl, N, N, qs, ls, us, [("solve",), ("get",)], path="stock", rho=rho0, **kw)[1]
    gb = R.Batch(mdl, N, Nmax, q, l, u, gpu_settings(**kw), rho=rho0)
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    _check(out, ref, n, m)
    assert (out["rho_updates"] == ref["rho_updates"]).all()
    assert out["rho_updates"].max() >= 1
    gb.free()
