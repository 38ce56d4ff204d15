"""Ruiz scaling (settings.scaling > 0, the OSQP default of 10) on the B200 path — SURVEY §8f-3.

The reference's recursive path is wrong with scaling on (it scales the global P / A but factors the unscaled
stage blocks), so the oracle is its STOCK path on the assembled P / A with the same settings
(F0: the same mathematics).  The kernels never form scaled data: they run the unscaled iteration with
sigma_i = sigma / (c D_i^2) and rho_j = rho_class(E_j l_j, E_j u_j) E_j^2 / c, which reproduces the
reference's scaled iteration to rounding.  Bar: D, E, c to 1e-12; status, iteration count, rho updates equal;
x, y, obj within 1e-6 relative.
"""
import ctypes as C
import os

import numpy as np
import pytest

import rldl_b200 as R
from rldl_b200 import mpc
from common import gpu_settings, ref_run, compare, scaled_problem, settings_dict

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
ADAPT = dict(adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5)


def _ref_out(g, tag, nmax, mmax):
    n, m = int(g["n"]), int(g["m"])
    return dict(status=g["status" + tag], iter=g["iter" + tag], obj=g["obj" + tag],
                x=np.pad(g["x" + tag], ((0, 0), (0, nmax - n))), y=np.pad(g["y" + tag], ((0, 0), (0, mmax - m))))


@pytest.mark.parametrize("name", ["C1", "C3", "C4"])
@pytest.mark.parametrize("mode", ["latency", "stream", "generic"])
def test_scaled_solves_match_golden(name, mode):
    """Every ADMM kernel on Ruiz-scaled problems: the BASELINE shapes have streaming / latency-mode variants whose stage
    records carry sigma per variable and the step size per row (Shape<..., RZ = 1>); the runtime-shape kernels serve
    everything else."""
    g = np.load(os.path.join(GOLD, f"scaled_{name}.npz"))
    mdl, N, x0, q, l, u = scaled_problem(name, int(g["batch"]))
    n, m = int(g["n"]), int(g["m"])

    def tune(gb):
        assert gb.select_kernel(2) == 0      # the shape-specialised kernels serve the scaled BASELINE shapes
        if mode == "generic":
            assert gb.select_kernel(1) == 0
        else:
            gb.set_tuning(lat_below=100000 if mode == "latency" else 0)

    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(scaling=10))
    tune(gb)
    D, E, c = gb.get_scaling()
    assert np.abs(D[:, :n] / g["D"] - 1).max() < 1e-12 and np.abs(E[:, :m] / g["E"] - 1).max() < 1e-12
    assert np.abs(c / g["c"] - 1).max() < 1e-12
    assert gb.solve() == 0
    bad = compare(gb.get(), _ref_out(g, "", q.shape[1], l.shape[1]), n, m)
    gb.free()
    assert not bad, bad[:3]
    ga = R.Batch(mdl, N, N, q, l, u, gpu_settings(scaling=10, **ADAPT), rho=g["rho_a"])
    tune(ga)
    assert ga.solve() == 0
    out = ga.get()
    ga.free()
    bad = compare(out, _ref_out(g, "_a", q.shape[1], l.shape[1]), n, m)
    assert not bad, bad[:3]
    assert (out["rho_updates"] == g["rho_updates_a"]).all()


def test_scaled_batch_on_streaming_kernels_matches_reference_sample():
    """C5-shaped batch with the OSQP default scaling = 10, large enough for the streaming kernel with pre-emption and the
    hand-off to the latency-mode kernel: every instance against the runtime-shape kernels, a sample against the
    reference's stock path."""
    mdl, N, x0, q, l, u = scaled_problem("C3", 6000)
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(scaling=10))
    assert gb.select_kernel(2) == 0
    gb.set_tuning(lat_below=512)
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    gb.free()
    gg = R.Batch(mdl, N, N, q[:400], l[:400], u[:400], gpu_settings(scaling=10))
    assert gg.select_kernel(1) == 0 and gg.solve() == 0
    bad = compare({k: v[:400] for k, v in out.items()}, gg.get(), n, m)
    gg.free()
    assert not bad, bad[:3]
    idx = np.sort(np.random.default_rng(8).choice(6000, 24, replace=False))
    ref = ref_run(mdl, N, N, q[idx], l[idx], u[idx], [("solve",), ("get",)], path="stock", scaling=10)[1]
    bad = compare({k: v[idx] for k, v in out.items()}, ref, n, m)
    assert not bad, bad[:3]


def test_scaled_updates_and_warm_start_match_reference():
    """The update entry points on a scaled workspace: new bounds are typed after scaling (osqp.c:815-817), a new
    linear cost keeps D, E, c (osqp.c:769-771), warm starts are given in unscaled variables (osqp.c:942-945),
    osqp_update_rho re-types and re-factorises."""
    mdl, N, x0, q, l, u = scaled_problem("C1", 8)
    rng = np.random.default_rng(3)
    n, m = N * 6, N * 10 + 4
    q2 = q + 0.3 * rng.standard_normal(q.shape)
    _, l2, u2 = mdl.vectors(mpc.sample_x0(mdl, 8, 77), N)
    l2[:, 10:14] = -1e30; u2[:, 10:14] = 1e30          # stage-1 state box loose: row types change
    xw, yw = 0.1 * rng.standard_normal(q.shape), 0.1 * rng.standard_normal(l.shape)
    rho2 = np.where(np.arange(8) % 2 == 0, 0.02, 1.5)
    ops = [("solve",), ("get",), ("update_lin_cost", q2), ("update_bounds", l2, u2), ("solve",), ("get",),
           ("warm_start", xw, yw), ("solve",), ("get",), ("update_rho", rho2), ("solve",), ("get",)]
    res = ref_run(mdl, N, N, q, l, u, ops, path="stock", scaling=10, warm_start=1)
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(scaling=10, warm_start=1))
    for i, op in enumerate(ops):
        if op[0] == "get":
            bad = compare(gb.get(), res[i], n, m)
            assert not bad, (i, bad[:3])
        else:
            assert getattr(gb, op[0])(*op[1:]) == 0, op[0]
    gb.free()


def test_default_settings_work_on_the_recursive_entry_points():
    """osqp_set_default_settings (scaling = 10, adaptive rho, eps 1e-3) + osqp_setup_recursive, untouched: the
    setup the reference documents.  Checked against the reference's stock path with the same settings and the
    automatic adaptive-rho interval both sides use without wall-clock input (4 x check_termination)."""
    from test_api_gpu import Scalar
    mdl, N, x0, q, l, u = scaled_problem("C3", 1, Nmax=51)
    s = R.default_settings()
    assert s.scaling == 10 and s.adaptive_rho == 1 and s.polish == 0
    s.verbose = 0
    sc = Scalar(mdl, N, 51, q[0], l[0], u[0], s)
    assert sc.rc == 0
    w = sc.w.contents
    assert w.scaling is not None
    assert sc.solve() == 0
    out = sc.out()
    qs, ls, us = mdl.vectors(x0, N)
    n, m = sc.dims()
    kw = {k: getattr(s, k) for k in ("eps_abs", "eps_rel", "rho", "sigma", "alpha", "max_iter", "scaling",
                                      "adaptive_rho", "adaptive_rho_tolerance", "warm_start")}
    kw["adaptive_rho_interval"] = 100
    qs[:, :n] = q[:, :n]
    ref = ref_run(mdl, N, N, qs, ls, us, [("solve",), ("get",)], path="stock", **kw)[1]
    bad = compare(out, ref, n, m)
    assert not bad, bad
    assert sc.info.rho_updates == ref["rho_updates"][0]
    sc.free()


def test_generic_qp_with_default_scaling():
    """osqp_setup / osqp_solve with scaling = 10 on the reference's basic_qp (tests/basic_qp:
    x = (0.3, 0.7), y = (-2.9, 0, 0.2, 0), obj 1.88) and, for status / iteration count, against the oracle's scaled
    restatement (pinned to the reference, tests/test_oracle_cpu.py)."""
    from test_api_gpu import _generic
    from oracle import rldl_oracle as O
    i64 = lambda *v: np.array(v, np.int64)
    P = (2, 2, i64(0, 1, 3), i64(0, 0, 1), np.array([4.0, 1.0, 2.0]))          # upper triangle of [[4,1],[1,2]]
    A = (4, 2, i64(0, 2, 5), i64(0, 1, 0, 2, 3), np.ones(5))
    Pd = np.array([[4.0, 1.0], [1.0, 2.0]]); Ad = np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0], [0.0, 1.0]])
    q = np.array([1.0, 1.0]); l = np.array([1.0, 0.0, 0.0, -1e30]); u = np.array([1.0, 0.7, 0.7, 1e30])
    kw = dict(scaling=10, eps_abs=1e-8, eps_rel=1e-8, adaptive_rho=1, adaptive_rho_interval=50)
    L, w, rc, keep = _generic(P, A, q, l, u, **kw)
    assert rc == 0 and L.osqp_solve(w) == 0
    info = w.contents.info.contents
    sol = w.contents.solution.contents
    x = np.ctypeslib.as_array(sol.x, (2,)); y = np.ctypeslib.as_array(sol.y, (4,))
    assert info.status_val == 1
    assert np.allclose(x, [0.3, 0.7], atol=1e-6) and abs(info.obj_val - 1.88) < 1e-6
    assert np.allclose(y, [-2.9, 0.0, 0.2, 0.0], atol=1e-5)
    st = O.Settings(**{k: v for k, v in settings_dict(**kw).items() if k in vars(O.Settings())})
    want = O.Solver(Pd, Ad, q, l, u, O.DenseKKT(Pd, Ad, st.sigma), st).solve()
    assert (info.iter, info.rho_updates) == (want["iter"], want["rho_updates"])
    assert np.abs(x - want["x"]).max() < 1e-6 and np.abs(y - want["y"]).max() < 1e-6
    assert L.osqp_cleanup(w) == 0
