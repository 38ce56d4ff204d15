"""Edge cases of the hot path through the C ABI, each against the unmodified reference
(oracle/_ref) on identical inputs: shortest horizons, Nmax > N, a batch of one, iteration limits
that do not fall on a termination check, no termination checks at all, non-default ADMM
parameters, and batches whose instances have DIFFERENT constraint types per row (loose /
equality / inequality, src/auxil.c:79-101), on every ADMM kernel.

Oracle choice: the reference's recursive path where it is sound, its stock path on the assembled
P/A otherwise (the two agree to 1e-13 on the same mathematics; F3/F4 and the
N = Nmax = 1 failure found here make the recursive path unusable for those inputs).
"""
import numpy as np
import pytest

import rldl_b200 as R
from rldl_b200 import mpc
from common import gpu_settings, ref_run, compare, problem

pytestmark = pytest.mark.gpu


def _gpu(mdl, N, Nmax, q, l, u, kernel=None, rho=None, **kw):
    gb = R.Batch(mdl, N, Nmax, q, l, u, gpu_settings(**kw), rho=rho)
    if kernel is not None:
        assert gb.select_kernel(kernel) == 0
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    gb.free()
    return out, n, m


def _c1(batch, N, Nmax=None, seed=1234):
    mdl = mpc.make_model(mpc.CONFIGS["C1"])
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, seed), N, Nmax)
    return mdl, q, l, u


@pytest.mark.parametrize("N,Nmax,path", [(1, 1, "stock"), (1, 2, "recursive"), (2, 2, "recursive"),
                                         (7, 20, "recursive"), (20, 21, "recursive")])
def test_short_horizons_and_nmax_above_n(N, Nmax, path):
    """N = 1 is a stage-0 block followed directly by the terminal block; Nmax > N leaves unused tail
    storage that must not leak into the solve.  (The reference's recursive path itself fails at
    N = Nmax = 1 — it runs to max_iter where its stock path needs 50 iterations — so that case is
    checked against the stock path.)"""
    mdl, q, l, u = _c1(6, N, Nmax)
    if path == "stock":
        qs, ls, us = mdl.vectors(mpc.sample_x0(mdl, 6), N, N)
        ref = ref_run(mdl, N, N, qs, ls, us, [("solve",), ("get",)], path="stock")[1]
    else:
        ref = ref_run(mdl, N, Nmax, q, l, u, [("solve",), ("get",)])[1]
    for kernel in (None, 0, 1):
        out, n, m = _gpu(mdl, N, Nmax, q, l, u, kernel)
        assert (n, m) == (N * 6, N * 10 + 4)
        bad = compare(out, ref, n, m)
        assert not bad, (kernel, bad[:3])


def test_batch_of_one():
    mdl, q, l, u = _c1(1, 20)
    ref = ref_run(mdl, 20, 20, q, l, u, [("solve",), ("get",)])[1]
    out, n, m = _gpu(mdl, 20, 20, q, l, u)
    assert not compare(out, ref, n, m)


@pytest.mark.parametrize("kw", [dict(max_iter=60), dict(max_iter=25), dict(max_iter=1),
                                dict(check_termination=0, max_iter=100),
                                dict(check_termination=7, max_iter=500),
                                dict(alpha=1.0, sigma=1e-3, rho=1.0),
                                dict(eps_abs=1e-3, eps_rel=1e-3, alpha=1.8),
                                dict(eps_abs=1e-9, eps_rel=0.0, max_iter=110)],
                         ids=["max_iter_60", "max_iter_25", "max_iter_1", "no_checks", "check_every_7",
                              "alpha1_sigma_rho", "loose_eps", "tight_eps_hits_limit"])
@pytest.mark.parametrize("kernel", [None, 0, 1], ids=["latency", "stream", "generic"])
def test_iteration_limits_and_parameters(kw, kernel):
    """Termination bookkeeping of src/osqp.c:354-627: the last iteration is checked even when it is
    not a multiple of check_termination, 'solved inaccurate' (status 2) is decided once after
    max_iter with 10x tolerances, check_termination = 0 disables the checks inside the loop."""
    mdl, q, l, u = _c1(24, 20)
    kw = dict(kw)
    rho = kw.pop("rho", None)                     # the step size goes in per instance (rldl_batch_setup's rho array)
    rho = None if rho is None else np.full(24, rho)
    ref = ref_run(mdl, 20, 20, q, l, u, [("solve",), ("get",)], rho=rho, **kw)[1]
    out, n, m = _gpu(mdl, 20, 20, q, l, u, kernel, rho=rho, **kw)
    bad = compare(out, ref, n, m)
    assert not bad, bad[:3]


@pytest.mark.parametrize("name,batch", [("C1", 48), ("C3", 24), ("C4", 12)])
@pytest.mark.parametrize("kernel", [None, 0, 1], ids=["latency", "stream", "generic"])
def test_per_instance_constraint_types_and_costs(name, batch, kernel):
    """Every instance gets its own mix of loose rows (|bound| >= 1e26 -> rho 1e-6), equality rows
    (u - l < 1e-4 -> 1e3 rho) and inequality rows, asymmetric bounds and a random linear cost, so the
    2-bit row classes of the stage records differ between instances.  Oracle = stock path (the
    recursive path reads the terminal rho from stage N-1's rows)."""
    cfg = mpc.CONFIGS[name]
    mdl = mpc.make_model(cfg)
    N, nx, nu = cfg["N"], mdl.nx, mdl.nu
    a, b = nx + nu, 2 * nx + nu
    q, l, u = mdl.vectors(0.3 * mpc.sample_x0(mdl, batch), N)
    rng = np.random.default_rng(99)
    q += 0.05 * rng.standard_normal(q.shape)
    for k in range(1, N):
        r = k * b
        loose = rng.random((batch, nx)) < 0.3                      # state rows: some unbounded
        l[:, r:r + nx][loose] = -1e30
        u[:, r:r + nx][loose] = 1e30
        onesided = rng.random((batch, nx)) < 0.2                   # some one-sided
        u[:, r:r + nx][onesided & ~loose] = 1e30
        eq = rng.random((batch, nu)) < 0.15                        # input rows: some pinned (equality)
        mid = 0.2 * (l[:, r + nx:r + a] + u[:, r + nx:r + a]) + 0.05 * (u[:, r + nx:r + a] - l[:, r + nx:r + a])
        l[:, r + nx:r + a][eq] = mid[eq]
        u[:, r + nx:r + a][eq] = mid[eq]
        shrink = rng.random((batch, nu)) < 0.3                     # some asymmetric
        u[:, r + nx:r + a][shrink & ~eq] *= 0.5
    ref = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",)], path="stock")[1]
    assert (ref["status"] == 1).mean() > 0.5                       # the case is not degenerate
    out, n, m = _gpu(mdl, N, N, q, l, u, kernel)
    bad = compare(out, ref, n, m)
    assert not bad, bad[:3]


def test_update_settings_between_solves():
    """Settings changed between solves (the osqp_update_* setters, include/osqp.h:258-419) take effect
    on the next solve: tolerances, max_iter, alpha, check_termination."""
    mdl, q, l, u = _c1(16, 20)
    kw1 = dict(eps_abs=1e-3, eps_rel=1e-3)
    kw2 = dict(eps_abs=1e-7, eps_rel=1e-7, alpha=1.4, check_termination=10, max_iter=900)
    r1 = ref_run(mdl, 20, 20, q, l, u, [("solve",), ("get",)], **kw1)[1]
    r2 = ref_run(mdl, 20, 20, q, l, u, [("solve",), ("get",)], **kw2)[1]
    gb = R.Batch(mdl, 20, 20, q, l, u, gpu_settings(**kw1))
    n, m, _, _ = gb.dims()
    assert gb.solve() == 0
    assert not compare(gb.get(), r1, n, m)
    assert gb.update_settings(gpu_settings(**kw2)) == 0
    assert gb.solve() == 0                                          # warm_start = 0: cold again
    bad = compare(gb.get(), r2, n, m)
    gb.free()
    assert not bad, bad[:3]


def test_non_convex_stage_cost_is_reported_like_the_reference():
    """A stage cost that leaves Q + sigma I + C' rho C indefinite: the reference's osqp_setup fails with
    OSQP_NONCVX_ERROR (src/osqp.c:214-216, "The problem seems to be non-convex"; its own tests/non_cvx).  The
    scalar setup returns that error; a batch keeps going, names the instances through
    rldl_batch_failed_factorisations and reports them as OSQP_NON_CVX (-7) with NaN solution at the solve."""
    import ctypes as C
    mdl, N, x0, q, l, u = problem("C1", 4, Nmax=21)
    mdl.Q = mdl.Q.copy()
    mdl.Q[1, 1] = -1.0e3                    # far below what sigma, rho and the dynamics rows' 1e3 rho add back
    gb = R.Batch(mdl, N, 21, q, l, u, gpu_settings())
    count, flag = gb.failed_factorisations()
    assert count == 4 and (flag > 0).all()   # 1 + stage of the first non-positive pivot
    assert gb.solve() == 0
    out = gb.get()
    assert (out["status"] == -7).all()
    n, m, _, _ = gb.dims()
    assert (out["x"][:, :n] == float(0x7fc00000)).all() or np.isnan(out["x"][:, :n]).all()
    gb.free()
    # the healthy model of the same shape factors cleanly
    mdl2, _, _, q2, l2, u2 = problem("C1", 4, Nmax=21)
    ok = R.Batch(mdl2, N, 21, q2, l2, u2, gpu_settings())
    assert ok.failed_factorisations()[0] == 0
    ok.free()
    # scalar API: setup refuses
    L = R.lib()
    sm = R.StageModel(mdl)
    d = sm.data
    q0, l0, u0 = (np.ascontiguousarray(v[0]) for v in (q, l, u))
    d.data.n, d.data.m = q.shape[1], l.shape[1]
    d.data.q = q0.ctypes.data_as(C.POINTER(C.c_double)); d.data.l = l0.ctypes.data_as(C.POINTER(C.c_double))
    d.data.u = u0.ctypes.data_as(C.POINTER(C.c_double))
    w = C.POINTER(R.OSQPWorkspace)()
    s = gpu_settings()
    assert L.osqp_setup_recursive(C.byref(w), C.byref(d), C.byref(s), 21, N, mdl.nx, mdl.nu, mdl.ny, mdl.nt) == 5


@pytest.mark.parametrize("kernel,below", [(0, 0), (0, 100000), (1, 0)], ids=["stream", "latency", "generic"])
def test_time_limit_is_enforced(kernel, below):
    """time_limit (src/osqp.c:389-405, :570-576): a solve that runs out of time stops with
    OSQP_TIME_LIMIT_REACHED (-6) — 'solved inaccurate' if the 10x tolerances hold — and returns its iterates;
    the kernels test the clock at the termination checks, so the iteration count is a multiple of
    check_termination, well below what the solve needs."""
    mdl, q, l, u = _c1(32, 20)
    full, n, m = _gpu(mdl, 20, 20, q, l, u, kernel, eps_abs=1e-9, eps_rel=1e-9, max_iter=4000)
    gb = R.Batch(mdl, 20, 20, q, l, u, gpu_settings(eps_abs=1e-9, eps_rel=1e-9, max_iter=4000, time_limit=1e-7))
    assert gb.select_kernel(kernel) == 0
    gb.set_tuning(lat_below=below)
    assert gb.solve() == 0
    out = gb.get()
    assert (out["status"] == -6).all(), out["status"]
    assert (out["iter"] == 25).all() and (full["iter"] > 25).all()
    assert np.isfinite(out["x"][:, :n]).all() and np.abs(out["x"][:, :n]).max() > 0
    # without the limit the same handle solves normally again
    assert gb.update_settings(gpu_settings(eps_abs=1e-9, eps_rel=1e-9, max_iter=4000, time_limit=0.0)) == 0
    assert gb.solve() == 0
    again = gb.get()
    gb.free()
    assert (again["iter"] == full["iter"]).all() and (again["status"] == full["status"]).all()


def test_hand_off_never_skips_unclaimed_instances():
    """with fewer resident lane groups than the hand-off threshold the streaming kernel used to stop
    claiming fresh instances once the unfinished count fell below it, leaving never-started instances with a
    stale status for the latency-mode kernel to skip.  One CTA per SM (1 184 groups at the C1 shape), threshold
    1 024, a warm-started batch larger than the resident groups whose instances all finish at their first check."""
    mdl, q, l, u = _c1(6000, 20)
    ref = R.Batch(mdl, 20, 20, q, l, u, gpu_settings(warm_start=1))
    assert ref.select_kernel(1) == 0
    assert ref.solve() == 0 and ref.solve() == 0     # second solve: warm-started at the optimum
    want = ref.get()
    ref.free()
    gb = R.Batch(mdl, 20, 20, q, l, u, gpu_settings(warm_start=1))
    gb.set_tuning(ctas_per_sm=1, lat_below=1024)
    assert gb.solve() == 0
    first = gb.get()
    assert gb.refactor() == 0                         # bench.py's step: refactor + solve without a status reset
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    gb.free()
    assert (out["status"] == 1).all()
    assert (out["iter"] == want["iter"]).all(), np.flatnonzero(out["iter"] != want["iter"])[:10]
    assert (first["iter"] > out["iter"]).any()
    assert np.abs(out["x"][:, :n] - want["x"][:, :n]).max() < 1e-6


def test_batch_setup_validates_bounds():
    """validate_data (src/auxil.c:794-887): l > u is refused at setup — on the device for a batch."""
    mdl, q, l, u = _c1(64, 20)
    l = l.copy()
    l[37, 11] = u[37, 11] + 1.0
    with pytest.raises(RuntimeError, match="code 1"):
        R.Batch(mdl, 20, 20, q, l, u, gpu_settings())
