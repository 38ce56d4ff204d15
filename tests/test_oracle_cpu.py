"""CPU suite: the numpy restatement (oracle/rldl_oracle.py) against the committed golden
vectors (generated from the unmodified reference by tests/golden/make_golden.py), against
the live reference when oracle/_ref is present, and against the reference's own
known-answer fixtures."""
import os

import numpy as np
import pytest

from common import problem, ref_run, settings_dict, rel_err, compare, REL_TOL
from oracle import rldl_oracle as O, ref_py, scenarios
from rldl_b200 import mpc

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def as_out(g, n, m, nmax, mmax):
    B = len(g["iter"])
    x, y = np.zeros((B, nmax)), np.zeros((B, mmax))
    x[:, :n], y[:, :m] = g["x"], g["y"]
    return dict(status=g["status"], iter=g["iter"], obj=g["obj"], x=x, y=y)


@pytest.mark.parametrize("name,take", [("C1", 8), ("C3", 2), ("C4", 2)])
def test_oracle_cold_solve_matches_golden(name, take):
    g = gold(f"cold_{name}")
    mdl, N, x0, q, l, u = problem(name, int(g["batch"]))
    ob = O.OracleBatch(mdl, N, N, q[:take], l[:take], u[:take], settings_dict())
    ob.solve()
    out = ob.get()
    n, m = ob.dims()
    ref = as_out(g, n, m, q.shape[1], l.shape[1])
    ref = {k: v[:take] for k, v in ref.items()}
    assert not compare(out, ref, n, m)
    for k in range(2):   # one KKT solve incl. the z_tilde step (qdldl_interface.c:559-585)
        assert rel_err(ob.kkt_solve(k, g["kkt_rhs"][k]), g["kkt_sol"][k]) < 1e-9


def test_oracle_adaptive_rho_matches_golden():
    g = gold("adaptive_C4")
    take = 2
    mdl, N, x0, q, l, u = problem("C4", int(g["batch"]))
    ob = O.OracleBatch(mdl, N, N, q[:take], l[:take], u[:take],
                       settings_dict(adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5),
                       rho=g["rho"][:take])
    ob.solve()
    out = ob.get()
    n, m = ob.dims()
    ref = {k: v[:take] for k, v in as_out(g, n, m, q.shape[1], l.shape[1]).items()}
    assert not compare(out, ref, n, m)
    assert (out["rho_updates"] == g["rho_updates"][:take]).all()


def test_oracle_receding_horizon_matches_golden():
    g = gold("receding_C1")
    cfg = mpc.CONFIGS["C1"]
    mdl, N = mpc.make_model(cfg), cfg["N"]
    B, T = int(g["batch"]), int(g["T"])
    x0 = mpc.sample_x0(mdl, B, int(g["seed"]))
    q, l, u = mdl.vectors(x0, N, N + 1)
    ob = O.OracleBatch(mdl, N, N + 1, q, l, u, settings_dict(warm_start=1))
    fb = [g[f"u0_{t}"] for t in range(T)]
    steps = scenarios.receding_horizon(ob, mdl, N, N + 1, x0, T, int(g["shrink_every"]), feedback=fb)
    for t, s in enumerate(steps):
        assert s["N"] == int(g[f"N_{t}"])
        assert (s["status"] == g[f"status_{t}"]).all(), t
        assert (s["iter"] == g[f"iter_{t}"]).all(), (t, s["iter"], g[f"iter_{t}"])
        assert rel_err(s["x"], g[f"x_{t}"]) < REL_TOL and rel_err(s["y"], g[f"y_{t}"]) < REL_TOL


def test_oracle_update_entry_points_match_golden():
    """update_lin_cost / partial_update_bounds / warm_start(x, y) / update_bounds + warm_start_x between
    warm-started solves, and per-instance update_rho on the dense path (fixture: make_golden.updates)."""
    import updates_script as US
    g = gold("updates_C1")
    mdl, N, q, l, u = US.inputs(g)
    take = 3
    US.replay(O.OracleBatch(mdl, N, N + 1, q[:take], l[:take], u[:take], settings_dict(warm_start=1)), g, take)
    n, m = int(g["n"]), int(g["m"])
    ob = O.OracleBatch(mdl, N, N, q[:take, :n], l[:take, :m], u[:take, :m], settings_dict(), path="stock")
    ob.solve(); ob.update_rho(g["rho2"][:take]); ob.solve()
    US.check(ob.get(), g, "rho", n, m, take)


def test_oracle_primal_infeasible_matches_golden():
    g = gold("infeasible_C1")
    mdl, N, x0, q, l, u = problem("C1", 4)
    b = mdl.nx + mdl.ny
    l[:, b:b + mdl.nx] = 10.0; u[:, b:b + mdl.nx] = 11.0
    ob = O.OracleBatch(mdl, N, N, q[:2], l[:2], u[:2], settings_dict())
    ob.solve()
    out = ob.get()
    assert (out["status"] == g["status"][:2]).all() and (out["iter"] == g["iter"][:2]).all()
    assert (out["x"][:, :ob.dims()[0]] == O.OSQP_NAN).all()


def test_oracle_basic_qp_known_answer():
    """reference tests/basic_qp/generate_problem.py:21-23: x=(0.3,0.7), y=(-2.9,0,0.2,0), obj 1.88."""
    P = np.array([[4., 1.], [1., 2.]]); q = np.ones(2)
    A = np.array([[1., 1.], [1., 0.], [0., 1.], [0., 1.]])
    l = np.array([1., 0., 0., -O.OSQP_INFTY]); u = np.array([1., 0.7, 0.7, O.OSQP_INFTY])
    s = O.Settings(scaling=0, polish=0, adaptive_rho=0, eps_abs=1e-8, eps_rel=1e-8, warm_start=0, max_iter=20000)
    sv = O.Solver(P, A, q, l, u, O.DenseKKT(P, A, s.sigma), s)
    out = sv.solve()
    assert out["status"] == O.SOLVED
    assert np.allclose(out["x"], [0.3, 0.7], atol=1e-5) and np.allclose(out["y"], [-2.9, 0, 0.2, 0], atol=1e-4)
    assert abs(out["obj"] - 1.88) < 1e-5
    # warm start from the optimum terminates at the first check (test_basic_qp.h:857-905, 1 iteration there
    # with check_termination=1)
    s.check_termination = 1
    sv.warm_start(out["x"], out["y"])
    assert sv.solve()["iter"] == 1


def test_oracle_qdldl_osqp_kkt_known_answer():
    """QDLDL's test_osqp_kkt.h:1-30 — a quasidefinite OSQP-shaped KKT system, solved through
    the same [P+sigma I, A'; A, -1/rho] structure the dense path factors."""
    rng = np.random.default_rng(3)
    n, m = 5, 4
    L = rng.standard_normal((n, n)); P = L @ L.T
    A = rng.standard_normal((m, n)); rho = np.array([0.1, 100.0, 0.1, 1e-6]); sigma = 1e-6
    k = O.DenseKKT(P, A, sigma).factor(rho)
    rhs = rng.standard_normal(n + m)
    K = np.block([[P + sigma * np.eye(n), A.T], [A, -np.diag(1 / rho)]])
    w = np.linalg.solve(K, rhs)
    want = w.copy(); want[n:] = rhs[n:] + w[n:] / rho
    assert rel_err(k.solve(rhs), want) < 1e-9


def test_stage_kkt_equals_dense_kkt_and_horizon_update():
    """F0 of SURVEY §8c restated: the stage recursion solves the same KKT system as the
    assembled one, also after growing / shrinking the horizon from the saved S_k."""
    cfg = mpc.CONFIGS["C1"]
    mdl = mpc.make_model(cfg)
    x0 = mpc.sample_x0(mdl, 1)
    q, l, u = mdl.vectors(x0, 20, 24)
    rho = 0.1
    sk = O.StageKKT(mdl.dense_blocks(), mdl.nx, mdl.nu, mdl.ny, mdl.nt, 20, 1e-6)
    rng = np.random.default_rng(0)
    for N in (20, 23, 18, 24):
        Pu, A = mdl.assemble(N)
        Pu = Pu.toarray(); P = Pu + Pu.T - np.diag(np.diag(Pu)); A = A.toarray()
        n, m = P.shape[0], A.shape[0]
        rv = O.set_rho_vec(l[0, :m], u[0, :m], rho)
        if N == 20:
            sk.factor(rv)
        else:
            sk.set_horizon(N, rv)
        rhs = rng.standard_normal(n + m)
        assert rel_err(sk.solve(rhs), O.DenseKKT(P, A, 1e-6).factor(rv).solve(rhs)) < 1e-9, N


@pytest.mark.skipif(not ref_py.available(), reason="oracle/_ref not built")
def test_oracle_matches_live_reference_both_paths():
    mdl, N, x0, q, l, u = problem("C1", 4, Nmax=21)
    for path in ("recursive", "stock"):
        ref = ref_run(mdl, N, N + 1, q, l, u, [("solve",), ("get",)], path=path)[1]
        ob = O.OracleBatch(mdl, N, N + 1, q, l, u, settings_dict(), path=path)
        ob.solve()
        n, m = ob.dims()
        assert not compare(ob.get(), ref, n, m), path


@pytest.mark.parametrize("name,take", [("C1", [0, 1, 5]), ("C3", [1])])
def test_oracle_ruiz_scaling_matches_golden(name, take):
    """settings.scaling = 10: the restated scale_data (src/scaling.c:44-156) gives the reference's D, E, c to the
    last bit, and the scaled ADMM loop with unscaled termination (src/auxil.c:251-515) the same status,
    iteration count, rho updates, x, y, obj — cold and with adaptive rho (golden: reference stock path)."""
    from common import scaled_problem
    g = gold(f"scaled_{name}")
    mdl, N, x0, q, l, u = scaled_problem(name, int(g["batch"]))
    n, m = int(g["n"]), int(g["m"])
    for tag, kw, rho in (("", dict(scaling=10), None),
                         ("_a", dict(scaling=10, adaptive_rho=1, adaptive_rho_interval=25, adaptive_rho_tolerance=1.5),
                          g["rho_a"])):
        ob = O.OracleBatch(mdl, N, N, q[take], l[take], u[take], settings_dict(**kw), path="stock",
                           rho=None if rho is None else rho[take])
        for j, k in enumerate(take):
            assert np.array_equal(ob.sv[j].D, g["D"][k]) and np.array_equal(ob.sv[j].E, g["E"][k])
            assert ob.sv[j].c == g["c"][k]
        ob.solve()
        out = ob.get()
        ref = dict(status=g["status" + tag][take], iter=g["iter" + tag][take], obj=g["obj" + tag][take],
                   x=np.pad(g["x" + tag][take], ((0, 0), (0, q.shape[1] - n))),
                   y=np.pad(g["y" + tag][take], ((0, 0), (0, l.shape[1] - m))))
        assert not compare(out, ref, n, m), tag
        if tag:
            assert (out["rho_updates"] == g["rho_updates_a"][take]).all()


@pytest.mark.parametrize("name,take", [("C1", list(range(16))), ("C3", [0, 1, 5])])
def test_oracle_polish_matches_golden(name, take):
    """polish = 1: the restated polish (src/polish.c:212-350 — active-set guess, reduced KKT with delta
    regularisation, iterative refinement, normal-cone projection, acceptance test) reproduces the reference's
    status_polish (accepted and rejected cases) and its polished x, y, obj."""
    import sys
    sys.path.insert(0, GOLD)
    from make_golden import POLISH
    g = gold(f"polish_{name}")
    batch, prob, kw = POLISH[name]
    mdl, N, x0, q, l, u = prob(name, batch)
    n, m = int(g["n"]), int(g["m"])
    ob = O.OracleBatch(mdl, N, N, q[take], l[take], u[take], settings_dict(**kw), path="stock")
    ob.solve()
    out = ob.get()
    assert (out["status_polish"] == g["status_polish"][take]).all()
    assert set(g["status_polish"].tolist()) >= {1, -1}
    ref = dict(status=g["status"][take], iter=g["iter"][take], obj=g["obj"][take],
               x=np.pad(g["x"][take], ((0, 0), (0, q.shape[1] - n))), y=np.pad(g["y"][take], ((0, 0), (0, l.shape[1] - m))))
    assert not compare(out, ref, n, m)
