"""GPU parity of the update / warm-start / matrix-update / combine entry points (SURVEY §8 rows that sit
either side of the solve): every call goes through the C ABI; expectations come from the committed golden
vectors of the unmodified reference (tests/golden/updates_C1.npz, cold_C1.npz) or from a fresh workspace
built on the already-updated problem."""
import ctypes as C
import os

import numpy as np
import pytest

import rldl_b200 as R
import updates_script as US
from common import gpu_settings, settings_dict, problem, compare, ref_run
from oracle import rldl_oracle as O
from rldl_b200 import mpc

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DP = C.POINTER(C.c_double)


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


class Scalar:
    """One workspace behind the reference's scalar entry points (osqp_setup_recursive & co)."""

    def __init__(self, mdl, N, Nmax, q, l, u, settings, combine=None):
        self.L, self.mdl, self.sm = R.lib(), mdl, R.StageModel(mdl)
        self.d = self.sm.data
        self.q, self.l, self.u = (np.ascontiguousarray(v, np.float64).copy() for v in (q, l, u))
        self.d.data.n, self.d.data.m = len(self.q), len(self.l)
        self.d.data.q, self.d.data.l, self.d.data.u = (v.ctypes.data_as(DP) for v in (self.q, self.l, self.u))
        self.s = settings
        self.w = C.POINTER(R.OSQPWorkspace)()
        dims = (mdl.nx, mdl.nu, mdl.ny, mdl.nt)
        if combine is None:
            self.rc = self.L.osqp_setup_recursive(C.byref(self.w), C.byref(self.d), C.byref(self.s), Nmax, N, *dims)
            self._solve = self.L.osqp_solve_recursive
        else:
            self.rc = self.L.osqp_setup_combine_recursive(C.byref(self.w), C.byref(self.d), C.byref(self.s), Nmax,
                                                          combine[0], combine[1], *dims)
            self._solve = self.L.osqp_solve_combine_recursive

    @property
    def info(self):
        return self.w.contents.info.contents

    def dims(self):
        return int(self.w.contents.data.contents.n), int(self.w.contents.data.contents.m)

    def solve(self):
        return self._solve(self.w)

    def out(self):
        n, m = self.dims()
        sol = self.w.contents.solution.contents
        x = np.ctypeslib.as_array(sol.x, (n,)).copy()[None]
        y = np.ctypeslib.as_array(sol.y, (m,)).copy()[None]
        i = self.info
        return dict(status=np.array([i.status_val]), iter=np.array([i.iter]), obj=np.array([i.obj_val]), x=x, y=y)

    def csc_values(self, which):
        M = getattr(self.w.contents.data.contents, which).contents
        nnz = int(M.p[M.n])
        return np.ctypeslib.as_array(M.x, (nnz,)).copy()

    def free(self):
        if self.w:
            assert self.L.osqp_cleanup(self.w) == 0
            self.w = None


def test_batch_update_entry_points_match_golden():
    """rldl_batch_update_lin_cost / _partial_update_bounds / _warm_start(x,y) / _update_bounds /
    _warm_start(x, NULL) between warm-started solves: the reference's sequence, instance by instance."""
    g = gold("updates_C1")
    mdl, N, q, l, u = US.inputs(g)
    B = int(g["batch"])
    gb = R.Batch(mdl, N, N + 1, q, l, u, gpu_settings(warm_start=1))
    US.replay(gb, g, B)
    gb.free()


def test_batch_update_rho_matches_golden():
    """rldl_batch_update_rho with one rho per instance, then a cold solve (reference: osqp_update_rho on its
    stock path, src/osqp.c:1268 — its recursive path segfaults there)."""
    g = gold("updates_C1")
    mdl, N, q, l, u = US.inputs(g)
    B, n, m = int(g["batch"]), int(g["n"]), int(g["m"])
    gb = R.Batch(mdl, N, N + 1, q, l, u, gpu_settings())
    assert gb.solve() == 0
    assert gb.update_rho(g["rho2"]) == 0
    assert gb.solve() == 0
    US.check(gb.get(), g, "rho", n, m, B)
    bad = g["rho2"].copy(); bad[1] = -1.0
    assert gb.update_rho(bad) == 1            # "rho must be positive" (osqp.c:1277)
    gb.free()


def test_scalar_update_entry_points_match_golden():
    """The same sequence on one workspace through osqp_update_lin_cost, osqp_partial_update_bounds,
    osqp_warm_start, osqp_update_lower_bound + osqp_update_upper_bound, osqp_warm_start_x."""
    g = gold("updates_C1")
    mdl, N, q, l, u = US.inputs(g)
    n, m, s, e = int(g["n"]), int(g["m"]), int(g["s"]), int(g["e"])
    k = 2                                                   # the instance with the most varied iteration counts
    ws = Scalar(mdl, N, N + 1, q[k], l[k], u[k], gpu_settings(warm_start=1))
    assert ws.rc == 0
    L, w = ws.L, ws.w
    gk = {key: (g[key][k:k + 1] if g[key].ndim and len(g[key]) == int(g["batch"]) else g[key]) for key in g.files}
    assert ws.solve() == 0
    US.check(ws.out(), gk, 0, n, m, 1)
    q2 = np.ascontiguousarray(g["q2"][k]); l2 = np.ascontiguousarray(g["l2"][k, s:e]); u2 = np.ascontiguousarray(g["u2"][k, s:e])
    assert L.osqp_update_lin_cost(w, q2.ctypes.data) == 0
    assert L.osqp_partial_update_bounds(w, s, e, l2.ctypes.data, u2.ctypes.data) == 0
    assert ws.solve() == 0
    US.check(ws.out(), gk, 1, n, m, 1)
    xw, yw = np.ascontiguousarray(g["xw"][k]), np.ascontiguousarray(g["yw"][k])
    assert L.osqp_warm_start(w, xw.ctypes.data, yw.ctypes.data) == 0
    assert ws.solve() == 0
    US.check(ws.out(), gk, 2, n, m, 1)
    lt, ut = np.ascontiguousarray(g["lt"][k]), np.ascontiguousarray(g["ut"][k])
    # lt <= old u and ut >= lt here, so the one-sided updates pass their own l <= u checks in either order
    lo_first = bool((lt[:m] <= ws.u[:m]).all())
    calls = [(L.osqp_update_lower_bound, lt), (L.osqp_update_upper_bound, ut)]
    for fn, v in (calls if lo_first else calls[::-1]):
        assert fn(w, v.ctypes.data) == 0
    assert L.osqp_warm_start_x(w, xw.ctypes.data) == 0
    assert ws.solve() == 0
    US.check(ws.out(), gk, 3, n, m, 1)
    # osqp_warm_start_y alone keeps x and z of the previous solve (osqp.c:976-994) — compared with a batch
    # of one that went through the same sequence
    assert L.osqp_warm_start_y(w, yw.ctypes.data) == 0
    assert ws.solve() == 0
    gb = R.Batch(mdl, N, N + 1, q[k:k + 1], l[k:k + 1], u[k:k + 1], gpu_settings(warm_start=1))
    US.replay(gb, gk, 1)
    assert gb.warm_start(None, g["yw"][k:k + 1]) == 0 and gb.solve() == 0
    assert not compare(ws.out(), gb.get(), n, m)
    gb.free()
    # one-sided updates that cross the other bound are refused and leave the problem untouched (osqp.c:835-925)
    r = int(np.flatnonzero((np.abs(lt[:m]) < 1e20) & (np.abs(ut[:m]) < 1e20))[0])
    bad = lt.copy(); bad[r] = ut[r] + 1.0
    assert L.osqp_update_lower_bound(w, bad.ctypes.data) == 1
    bad = ut.copy(); bad[r] = lt[r] - 1.0
    assert L.osqp_update_upper_bound(w, bad.ctypes.data) == 1
    assert ws.solve() == 0 and ws.info.status_val == 1
    ws.free()


def test_update_P_A_on_recursive_workspace_matches_fresh_setup():
    """osqp_update_P / osqp_update_A / osqp_update_P_A (src/osqp.c:1016-1266) on an MPC workspace: the new
    global values are folded back into the stage blocks and the recursion is re-run; the result must equal a
    workspace set up on the new model.  A change that breaks the stage structure is refused."""
    N, Nmax = 12, 12
    m1, m2 = mpc.random_model(4, 2, seed=11), mpc.random_model(4, 2, seed=12)
    x0 = mpc.sample_x0(m1, 1, 5)
    q, l, u = m1.vectors(x0, N, Nmax)
    q2, l2, u2 = m2.vectors(x0, N, Nmax)
    a, b = Scalar(m1, N, Nmax, q[0], l[0], u[0], gpu_settings()), Scalar(m2, N, Nmax, q2[0], l2[0], u2[0], gpu_settings())
    assert a.rc == 0 and b.rc == 0
    assert b.solve() == 0
    want = b.out()
    n, m = b.dims()
    Px, Ax = b.csc_values("P"), b.csc_values("A")
    Px1, Ax1 = a.csc_values("P"), a.csc_values("A")
    assert Px.shape == Px1.shape and Ax.shape == Ax1.shape and not np.allclose(Ax, Ax1)
    L = a.L
    # (1) both at once, whole arrays (idx = NULL)
    assert L.osqp_update_P_A(a.w, Px.ctypes.data, None, len(Px), Ax.ctypes.data, None, len(Ax)) == 0
    assert L.osqp_update_lin_cost(a.w, np.ascontiguousarray(q2[0]).ctypes.data) == 0
    assert L.osqp_update_bounds(a.w, np.ascontiguousarray(l2[0]).ctypes.data, np.ascontiguousarray(u2[0]).ctypes.data) == 0
    assert a.solve() == 0
    assert not compare(a.out(), want, n, m)
    # (2) back to model 1 through the separate calls with explicit index lists
    idxP = np.arange(len(Px1), dtype=np.int64)
    idxA = np.arange(len(Ax1), dtype=np.int64)
    assert L.osqp_update_P(a.w, Px1.ctypes.data, idxP.ctypes.data, len(Px1)) == 0
    assert L.osqp_update_A(a.w, Ax1.ctypes.data, idxA.ctypes.data, len(Ax1)) == 0
    assert L.osqp_update_lin_cost(a.w, np.ascontiguousarray(q[0]).ctypes.data) == 0
    assert L.osqp_update_bounds(a.w, np.ascontiguousarray(l[0]).ctypes.data, np.ascontiguousarray(u[0]).ctypes.data) == 0
    L.osqp_update_warm_start(a.w, 0)
    assert a.solve() == 0
    fresh = Scalar(m1, N, Nmax, q[0], l[0], u[0], gpu_settings())
    assert fresh.solve() == 0
    assert not compare(a.out(), fresh.out(), n, m)
    # (3) an index list longer than the matrix: refused (osqp.c:1020-1031, 1098-1109)
    longP, longA = np.zeros(len(Px1) + 1, np.int64), np.zeros(len(Ax1) + 1, np.int64)
    assert L.osqp_update_P(a.w, np.zeros(len(longP)).ctypes.data, longP.ctypes.data, len(longP)) == 1
    assert L.osqp_update_A(a.w, np.zeros(len(longA)).ctypes.data, longA.ctypes.data, len(longA)) == 1
    # (4) one entry of one interior stage changed: the stages no longer share blocks -> refused
    one = np.array([Ax1[len(Ax1) // 2] + 0.25]); at = np.array([len(Ax1) // 2], np.int64)
    assert L.osqp_update_A(a.w, one.ctypes.data, at.ctypes.data, 1) == 1
    for ws in (a, b, fresh):
        ws.free()


def test_combine_variant_matches_golden_and_horizon_changes():
    """osqp_setup_combine_recursive (head N=15 + tail Nplus=5) solves the N=20 problem of cold_C1;
    osqp_update_Z_horizon grows / shrinks the tail and refuses to cut into the head
    (src/recursive_ldl.c:2359, 2761-2860)."""
    g = gold("cold_C1")
    mdl, N, x0, q, l, u = problem("C1", 1, Nmax=22)
    ws = Scalar(mdl, 20, 22, q[0], l[0], u[0], gpu_settings(), combine=(15, 5))
    assert ws.rc == 0 and ws.solve() == 0
    n, m = int(g["n"]), int(g["m"])
    assert ws.dims() == (n, m)
    ref = dict(status=g["status"][:1], iter=g["iter"][:1], obj=g["obj"][:1], x=g["x"][:1], y=g["y"][:1])
    assert not compare(ws.out(), ref, n, m)
    L = ws.L
    L.osqp_update_warm_start(ws.w, 0)
    Ncur = 20
    for dN in (+2, -5):
        assert L.osqp_update_Z_horizon(ws.w, C.byref(ws.d), dN) == 0
        Ncur += dN
        assert ws.solve() == 0
        qf, lf, uf = mdl.vectors(x0[:1], Ncur, 22)
        fresh = R.Batch(mdl, Ncur, 22, qf, lf, uf, gpu_settings())
        assert fresh.solve() == 0
        nn, mm, _, _ = fresh.dims()
        assert ws.dims() == (nn, mm)
        assert not compare(ws.out(), fresh.get(), nn, mm)
        fresh.free()
    assert L.osqp_update_Z_horizon(ws.w, C.byref(ws.d), -3) == -1      # 17 - 3 < head of 15
    assert L.osqp_update_Z_horizon(ws.w, C.byref(ws.d), +6) == -1      # 17 + 6 > Nmax
    assert L.osqp_update_X_horizon(ws.w, C.byref(ws.d), 1) == 0        # empty body in the reference
    assert ws.solve() == 0 and ws.dims() == (nn, mm)
    ws.free()


def oracle_settings():
    return O.Settings(**{k: v for k, v in settings_dict().items() if k in vars(O.Settings())})


def _generic(P, A, q, l, u, **kw):
    L = R.lib()
    Pc, Ac = R.make_csc(P), R.make_csc(A)
    d = R.OSQPData()
    d.n, d.m, d.P, d.A = P[1], A[0], C.pointer(Pc), C.pointer(Ac)
    d.q, d.l, d.u = (v.ctypes.data_as(DP) for v in (q, l, u))
    s = gpu_settings(**kw)
    w = C.POINTER(R.OSQPWorkspace)()
    rc = L.osqp_setup(C.byref(w), C.byref(d), C.byref(s))
    return L, w, rc, (Pc, Ac, d, s)


def test_generic_qp_dual_infeasible_and_primal_infeasible_status():
    """Known answers of the reference's own fixtures: tests/primal_dual_infeasibility (status codes
    OSQP_DUAL_INFEASIBLE = -4 with obj = -OSQP_INFTY, OSQP_PRIMAL_INFEASIBLE = -3 with obj = +OSQP_INFTY;
    src/auxil.c:606-643), iteration count against the oracle's dense path."""
    i64 = lambda *v: np.array(v, np.int64)
    # min x1 - x2  s.t. x >= 0, P = 0: unbounded below along (0, 1)
    P = (2, 2, i64(0, 1, 2), i64(0, 1), np.zeros(2))
    A = (2, 2, i64(0, 1, 2), i64(0, 1), np.ones(2))
    q = np.array([1.0, -1.0]); l = np.array([0.0, 0.0]); u = np.array([1e30, 1e30])
    L, w, rc, keep = _generic(P, A, q, l, u)
    assert rc == 0 and L.osqp_solve(w) == 0
    info = w.contents.info.contents
    sv = O.Solver(np.zeros((2, 2)), np.eye(2), q, l, u, O.DenseKKT(np.zeros((2, 2)), np.eye(2), 1e-6),
                  oracle_settings())
    want = sv.solve()
    assert want["status"] == -4
    assert (info.status_val, info.iter) == (-4, want["iter"])
    assert info.status == b"dual infeasible" and info.obj_val == -1e30
    assert L.osqp_cleanup(w) == 0
    # x1 <= -1 and x1 >= 1 at once through two rows
    A = (2, 2, i64(0, 2, 2), i64(0, 1), np.ones(2))
    P = (2, 2, i64(0, 1, 2), i64(0, 1), np.ones(2))
    l = np.array([1.0, -1e30]); u = np.array([1e30, -1.0]); q = np.zeros(2)
    L, w, rc, keep = _generic(P, A, q, l, u)
    assert rc == 0 and L.osqp_solve(w) == 0
    info = w.contents.info.contents
    Ad = np.array([[1.0, 0.0], [1.0, 0.0]])
    sv = O.Solver(np.eye(2), Ad, q, l, u, O.DenseKKT(np.eye(2), Ad, 1e-6), oracle_settings())
    want = sv.solve()
    assert want["status"] == -3
    assert (info.status_val, info.iter) == (-3, want["iter"])
    assert info.status == b"primal infeasible" and info.obj_val == 1e30
    assert L.osqp_cleanup(w) == 0


def test_setup_validation_errors():
    """validate_data / validate_settings refusals (src/auxil.c:794-1020): non-upper-triangular P, l > u,
    a bad setting; the workspace pointer stays NULL."""
    i64 = lambda *v: np.array(v, np.int64)
    A = (1, 2, i64(0, 1, 2), i64(0, 0), np.ones(2))
    q = np.ones(2); l = np.array([0.0]); u = np.array([1.0])
    Plow = (2, 2, i64(0, 2, 3), i64(0, 1, 1), np.array([1.0, 0.5, 1.0]))     # entry (1,0): below the diagonal
    L, w, rc, _ = _generic(Plow, A, q, l, u)
    assert rc != 0 and not w
    P = (2, 2, i64(0, 1, 2), i64(0, 1), np.ones(2))
    L, w, rc, _ = _generic(P, A, q, np.array([2.0]), u)
    assert rc != 0 and not w
    L, w, rc, _ = _generic(P, A, q, l, u, alpha=2.5)
    assert rc != 0 and not w
    L, w, rc, _ = _generic(P, A, q, l, u)
    assert rc == 0 and L.osqp_solve(w) == 0 and w.contents.info.contents.status_val == 1
    # setters validate like the reference (src/osqp.c:1321-1560)
    assert L.osqp_update_alpha(w, 2.0) == 1 and L.osqp_update_alpha(w, 1.5) == 0
    assert L.osqp_update_max_iter(w, 0) == 1 and L.osqp_update_max_iter(w, 10) == 0
    assert L.osqp_update_eps_abs(w, -1.0) == 1 and L.osqp_update_eps_rel(w, -1.0) == 1
    assert L.osqp_update_check_termination(w, -1) == 1
    assert L.osqp_update_polish(w, 2) == 1 and L.osqp_update_polish(w, 1) == 0 and L.osqp_update_polish(w, 0) == 0
    # max_iter = 10 with check_termination = 25: stops at 10 with "maximum iterations reached" (-2)
    L.osqp_update_warm_start(w, 0)
    assert L.osqp_update_eps_abs(w, 1e-14) == 0 and L.osqp_update_eps_rel(w, 1e-14) == 0
    assert L.osqp_solve(w) == 0
    info = w.contents.info.contents
    assert (info.status_val, info.iter) == (-2, 10)
    assert L.osqp_cleanup(w) == 0


def test_batch_update_model_matches_fresh_setup():
    """rldl_batch_update_model — the batched osqp_update_P_A (src/osqp.c:1166-1266): new stage-matrix values
    for a whole batch, full recursion re-run; equals a batch set up on the new model, cold and warm-started,
    on the specialised kernels (batch above and below the latency-mode hand-off) and the generic one."""
    N, Nmax = 12, 13
    m1, m2 = mpc.random_model(4, 2, seed=21), mpc.random_model(4, 2, seed=22)
    for batch, generic in ((24, False), (1500, False), (24, True)):
        x0 = mpc.sample_x0(m1, batch, 7)
        q, l, u = m1.vectors(x0, N, Nmax)
        q2, l2, u2 = m2.vectors(x0, N, Nmax)
        a = R.Batch(m1, N, Nmax, q, l, u, gpu_settings())
        b = R.Batch(m2, N, Nmax, q2, l2, u2, gpu_settings())
        if generic:
            assert a.select_kernel(1) == 0 and b.select_kernel(1) == 0
        assert a.solve() == 0                       # a has iterates and a factor of model 1
        assert a.update_model(m2) == 0
        assert (a.get()["status"] == -10).all()     # info reset like the reference's update (OSQP_UNSOLVED)
        assert a.update_lin_cost(q2) == 0 and a.update_bounds(l2, u2) == 0
        assert a.solve() == 0 and b.solve() == 0
        n, m, _, _ = b.dims()
        assert not compare(a.get(), b.get(), n, m)
        # a model of another shape is refused and leaves the handle usable
        assert a.update_model(mpc.random_model(3, 2, seed=5)) != 0
        assert a.solve() == 0
        assert not compare(a.get(), b.get(), n, m)
        a.free(); b.free()


def test_model_updates_and_combine_shrink_pinned_to_the_reference():
    """The sub-paths round 1 only checked against the product's own fresh setup, now against the UNMODIFIED
    reference's stock path (the same mathematics; the recursive workspace of the reference has no
    KKT to update, F1, and its combine-variant tail shrink diverges, F8):
      (1) osqp_update_P_A on a recursive workspace  vs  the reference's own osqp_update_P_A (src/osqp.c:1166-1266)
          with the same value arrays on a stock workspace, then osqp_update_lin_cost / osqp_update_bounds;
      (2) rldl_batch_update_model (the batched form)  vs  the same reference run;
      (3) osqp_update_Z_horizon(-2) after osqp_setup_combine_recursive(15 + 5)  vs  a reference stock setup at
          N = 18."""
    N, batch = 12, 6
    m1, m2 = mpc.random_model(4, 2, seed=31), mpc.random_model(4, 2, seed=32)
    x0 = mpc.sample_x0(m1, batch, 9)
    q, l, u = m1.vectors(x0, N)
    q2, l2, u2 = m2.vectors(x0, N)
    q2 = q2 + 0.1 * np.random.default_rng(2).standard_normal(q2.shape)
    Px2, Ax2 = ref_run(m2, N, N, q2, l2, u2, [("get_PA",)], path="stock")[0]
    ops = [("solve",), ("update_PA", Px2, Ax2), ("update_lin_cost", q2), ("update_bounds", l2, u2), ("solve",), ("get",)]
    res = ref_run(m1, N, N, q, l, u, ops, path="stock")
    assert res[1] == 0
    want = res[5]
    n, m = N * 6, N * 10 + 4
    # (1) scalar entry point, instance by instance
    for k in range(2):
        a = Scalar(m1, N, N, q[k], l[k], u[k], gpu_settings())
        assert a.rc == 0 and a.solve() == 0
        assert a.csc_values("P").shape == Px2.shape and a.csc_values("A").shape == Ax2.shape
        assert a.L.osqp_update_P_A(a.w, Px2.ctypes.data, None, len(Px2), Ax2.ctypes.data, None, len(Ax2)) == 0
        assert a.L.osqp_update_lin_cost(a.w, np.ascontiguousarray(q2[k]).ctypes.data) == 0
        assert a.L.osqp_update_bounds(a.w, np.ascontiguousarray(l2[k]).ctypes.data,
                                      np.ascontiguousarray(u2[k]).ctypes.data) == 0
        assert a.solve() == 0
        bad = compare(a.out(), {key: v[k:k + 1] for key, v in want.items()}, n, m)
        assert not bad, (k, bad)
        a.free()
    # (2) batched
    gb = R.Batch(m1, N, N, q, l, u, gpu_settings())
    assert gb.solve() == 0 and gb.update_model(m2) == 0
    assert gb.update_lin_cost(q2) == 0 and gb.update_bounds(l2, u2) == 0 and gb.solve() == 0
    bad = compare(gb.get(), want, n, m)
    gb.free()
    assert not bad, bad[:3]
    # (3) combine variant: tail shrink
    mdl, Nc, x0c, qc, lc, uc = problem("C1", 1, Nmax=22)
    ws = Scalar(mdl, 20, 22, qc[0], lc[0], uc[0], gpu_settings(), combine=(15, 5))
    assert ws.rc == 0 and ws.solve() == 0
    ws.L.osqp_update_warm_start(ws.w, 0)
    assert ws.L.osqp_update_Z_horizon(ws.w, C.byref(ws.d), -2) == 0 and ws.solve() == 0
    q18, l18, u18 = mdl.vectors(x0c[:1], 18)
    ref18 = ref_run(mdl, 18, 18, q18, l18, u18, [("solve",), ("get",), ("dims",)], path="stock")
    assert ws.dims() == tuple(ref18[2])
    bad = compare(ws.out(), ref18[1], *ref18[2])
    ws.free()
    assert not bad, bad
