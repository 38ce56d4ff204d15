"""oracle/rldl_oracle.py — CPU restatement (numpy) of the reference's recursive-LDL KKT path
and of the OSQP ADMM loop around it.  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product path
(librldl_b200.so has no CPU fallback).

Parity status: PINNED.  tests/test_oracle_cpu.py checks this restatement against
  * the UNMODIFIED reference compiled into oracle/_ref (live, when present), and
  * the committed vectors in tests/golden/ that were generated from oracle/_ref by
    tests/golden/make_golden.py,
on status, iteration count, x, y and objective, and against the reference's own
known-answer fixtures (tests/basic_qp: x=(0.3,0.7), obj 1.88; QDLDL's test_osqp_kkt).

What is restated (file:line into /root/reference):
  StageKKT.factor ....... LDL_factorize_recursive, src/recursive_ldl.c:1139-1318 (block form:
                          X_0 = Q_0+sigma I; S_k = diag(1/rho_k) + A_k X_k^-1 A_k';
                          X_{k+1} = Q_{k+1}+sigma I + E' S_k^-1 E), restart from a saved S_k:
                          LDL_update_from_pivot :946-1110
  StageKKT.solve ........ solve_linsys_qdldl, lin_sys/direct/qdldl/qdldl_interface.c:559-585
                          (block forward / diagonal / backward substitution, then
                          z_tilde = rhs_z + rho^-1 nu)
  set_rho_vec ........... src/auxil.c:79-101
  admm_solve ............ osqp_solve, src/osqp.c:288-641 with update_xz_tilde/update_x/
                          update_z/update_y src/auxil.c:164-228, project src/proj.c:4-14,
                          residuals/tolerances src/auxil.c:243-362, infeasibility certificates
                          :364-515, check_termination :684-789, compute_rho_estimate/adapt_rho
                          :13-77, osqp_update_rho src/osqp.c:1268-1319
"""
from __future__ import annotations

import numpy as np
import scipy.linalg as sla

OSQP_INFTY = 1e30
MIN_SCALING = 1e-4
RHO_MIN, RHO_MAX, RHO_EQ_OVER_RHO_INEQ, RHO_TOL = 1e-6, 1e6, 1e3, 1e-4
OSQP_NAN = float(0x7fc00000)  # include/constants.h:95 (sic: a finite number)

SOLVED, SOLVED_INACC, MAX_ITER, PRIM_INF, PRIM_INF_INACC, DUAL_INF, DUAL_INF_INACC, NON_CVX, UNSOLVED = \
    1, 2, -2, -3, 3, -4, 4, -7, -10


def set_rho_vec(l, u, rho):
    """src/auxil.c:79-101 — constraint typing: loose / equality / inequality."""
    rho = min(max(rho, RHO_MIN), RHO_MAX)
    rv = np.full(l.shape, rho)
    rv[(l < -OSQP_INFTY * MIN_SCALING) & (u > OSQP_INFTY * MIN_SCALING)] = RHO_MIN
    rv[(u - l) < RHO_TOL] = RHO_EQ_OVER_RHO_INEQ * rho
    return rv


class StageKKT:
    """Stage-wise block factorisation of the MPC KKT matrix in the reference's block order
    v_0, c_0, v_1, c_1, ..., v_N, c_N  (src/recursive_ldl.c:1153-1234)."""

    def __init__(self, blocks, nx, nu, ny, nt, N, sigma):
        self.bl, self.nx, self.nu, self.ny, self.nt, self.N, self.sigma = blocks, nx, nu, ny, nt, N, sigma
        self.a, self.b = nx + nu, nx + ny
        self.S = {}   # saved Schur complements S_k (the reference's X_even[k])

    # --- block accessors -------------------------------------------------
    def Qk(self, k):
        return self.bl["Q0"] if k == 0 else (self.bl["QN"] if k == self.N else self.bl["Qi"])

    def Ak(self, k):
        return self.bl["A0"] if k == 0 else (self.bl["AN"] if k == self.N else self.bl["Ai"])

    def vdim(self, k):
        return self.nu if k == 0 else (self.nx if k == self.N else self.a)

    def cdim(self, k):
        return self.nt if k == self.N else self.b

    def voff(self, k):
        return 0 if k == 0 else self.nu + (k - 1) * self.a

    def coff(self, k):
        return k * self.b

    @property
    def n(self):
        return self.N * self.a

    @property
    def m(self):
        return self.N * self.b + self.nt

    def factor(self, rho_vec, k0=0):
        """Run the recursion from stage k0 (k0>0 restarts from the saved S_{k0-1})."""
        nx, ny, N = self.nx, self.ny, self.N
        self.rho_vec = np.array(rho_vec[: self.m], float)
        if k0 == 0:
            self.X, self.Xc, self.Sc = {}, {}, {}
        X = self.Qk(k0) + self.sigma * np.eye(self.vdim(k0))
        if k0 > 0:
            Sinv = sla.cho_solve(self.Sc[k0 - 1], np.eye(self.b))
            X = X.copy()
            X[:nx, :nx] += Sinv[ny:, ny:]
        for k in range(k0, N + 1):
            self.X[k] = X
            self.Xc[k] = sla.cho_factor(X)   # raises LinAlgError on a non-positive pivot
            A = self.Ak(k)
            r = self.rho_vec[self.coff(k): self.coff(k) + self.cdim(k)]
            S = np.diag(1.0 / r) + A @ sla.cho_solve(self.Xc[k], A.T)
            self.S[k] = S
            self.Sc[k] = sla.cho_factor(S)
            if k < N:
                Sinv = sla.cho_solve(self.Sc[k], np.eye(self.b))
                X = self.Qk(k + 1) + self.sigma * np.eye(self.vdim(k + 1))
                X[:nx, :nx] += Sinv[ny:, ny:]          # E' S^-1 E on the state corner (:1217-1222)
        return self

    def set_horizon(self, Nnew, rho_vec):
        """LDL_update_from_pivot: restart at stage min(N,Nnew)-1's saved S (:963-997)."""
        k0 = min(self.N, Nnew)
        self.N = Nnew
        return self.factor(rho_vec, k0)

    def solve(self, rhs):
        """[x_tilde; z_tilde] for rhs = [rhs_x; rhs_z]  (qdldl_interface.c:559-585)."""
        nx, ny, N, n = self.nx, self.ny, self.N, self.n
        rv = [rhs[self.voff(k): self.voff(k) + self.vdim(k)].copy() for k in range(N + 1)]
        rc = [rhs[n + self.coff(k): n + self.coff(k) + self.cdim(k)].copy() for k in range(N + 1)]
        # forward
        for k in range(N + 1):
            if k > 0:   # K[v_k, c_{k-1}] = Aij' = -E'  => t_v -= Aij' (-S)^-1 t_c = -(S^-1 t_c)_dyn
                rv[k][:nx] -= sla.cho_solve(self.Sc[k - 1], rc[k - 1])[ny:]
            rc[k] -= self.Ak(k) @ sla.cho_solve(self.Xc[k], rv[k])
        # diagonal + backward
        wv, wc = [None] * (N + 1), [None] * (N + 1)
        for k in range(N, -1, -1):
            t = rc[k].copy()
            if k < N:   # K[c_k, v_{k+1}] = Aij = [0; -I 0]
                t[ny:] += wv[k + 1][:nx]
            wc[k] = -sla.cho_solve(self.Sc[k], t)
            wv[k] = sla.cho_solve(self.Xc[k], rv[k] - self.Ak(k).T @ wc[k])
        out = np.empty(n + self.m)
        for k in range(N + 1):
            out[self.voff(k): self.voff(k) + self.vdim(k)] = wv[k]
            c0 = self.coff(k)
            out[n + c0: n + c0 + self.cdim(k)] = rhs[n + c0: n + c0 + self.cdim(k)] + wc[k] / self.rho_vec[c0: c0 + self.cdim(k)]
        return out


class DenseKKT:
    """Generic (non-staged) KKT solve for the osqp_setup path: quasi-definite LDL' of
    [P+sigma I, A'; A, -diag(1/rho)]  (src/kkt.c:6, qdldl_interface.c:559)."""

    def __init__(self, P, A, sigma):
        self.P, self.A, self.sigma = P, A, sigma
        self.n, self.m = P.shape[0], A.shape[0]

    def factor(self, rho_vec, k0=0):
        self.rho_vec = np.array(rho_vec[: self.m], float)
        K = np.block([[self.P + self.sigma * np.eye(self.n), self.A.T],
                      [self.A, -np.diag(1.0 / self.rho_vec)]]) if self.m else self.P + self.sigma * np.eye(self.n)
        self.lu = sla.lu_factor(K)
        return self

    def solve(self, rhs):
        w = sla.lu_solve(self.lu, rhs)
        out = w.copy()
        out[self.n:] = rhs[self.n:] + w[self.n:] / self.rho_vec
        return out


# ---------------------------------------------------------------------------
# ADMM loop
# ---------------------------------------------------------------------------
def _inf(v):
    return float(np.max(np.abs(v))) if v.size else 0.0


class Settings:
    """The OSQPSettings fields the path reads (include/types.h:139-176, defaults constants.h:59-114)."""

    def __init__(self, **kw):
        self.rho, self.sigma, self.alpha = 0.1, 1e-6, 1.6
        self.max_iter, self.eps_abs, self.eps_rel = 4000, 1e-3, 1e-3
        self.eps_prim_inf, self.eps_dual_inf = 1e-4, 1e-4
        self.check_termination, self.warm_start = 25, 1
        self.adaptive_rho, self.adaptive_rho_interval, self.adaptive_rho_tolerance = 1, 0, 5.0
        self.scaling, self.polish = 10, 0
        for k, v in kw.items():
            setattr(self, k, v)


class Solver:
    """One QP instance: min 1/2 x'Px + q'x  s.t. l <= Ax <= u, with a pluggable KKT object
    (StageKKT for the recursive path, DenseKKT for the generic path)."""

    def __init__(self, P, A, q, l, u, kkt, settings: Settings):
        assert settings.scaling == 0 and settings.polish == 0, "out of scope (SURVEY §8b preconditions)"
        self.P, self.A = np.asarray(P, float), np.asarray(A, float)   # P full symmetric
        self.q, self.l, self.u = np.array(q, float), np.array(l, float), np.array(u, float)
        self.s, self.kkt = settings, kkt
        self.n, self.m = self.P.shape[0], self.A.shape[0]
        self.rho = min(max(settings.rho, RHO_MIN), RHO_MAX)
        self.rho_vec = set_rho_vec(self.l, self.u, self.rho)
        self.kkt.factor(self.rho_vec)
        self.x, self.z, self.y = np.zeros(self.n), np.zeros(self.m), np.zeros(self.m)
        self.rho_updates = 0

    # osqp_warm_start (src/osqp.c:929-994)
    def warm_start(self, x=None, y=None):
        self.s.warm_start = 1
        if x is not None:
            self.x = np.array(x[: self.n], float)
            self.z = self.A @ self.x
        if y is not None:
            self.y = np.array(y[: self.m], float)

    # osqp_update_bounds (src/osqp.c:784) -> update_rho_vec (src/auxil.c:103-145)
    def update_bounds(self, l, u):
        self.l, self.u = np.array(l[: self.m], float), np.array(u[: self.m], float)
        new = set_rho_vec(self.l, self.u, self.rho)
        if np.any(new != self.rho_vec):
            self.rho_vec = new
            self.kkt.factor(self.rho_vec)

    def update_lin_cost(self, q):
        self.q = np.array(q[: self.n], float)

    # osqp_update_rho (src/osqp.c:1268-1319)
    def update_rho(self, rho_new):
        self.rho = min(max(rho_new, RHO_MIN), RHO_MAX)
        self.rho_vec = set_rho_vec(self.l, self.u, self.rho)
        self.kkt.factor(self.rho_vec)

    def _residuals(self):
        Ax, Px, Aty = self.A @ self.x, self.P @ self.x, self.A.T @ self.y
        R = dict(pri=_inf(Ax - self.z), nz=_inf(self.z), nAx=_inf(Ax), dua=_inf(self.q + Px + Aty),
                 nq=_inf(self.q), nAty=_inf(Aty), nPx=_inf(Px))
        R["obj"] = 0.5 * float(self.x @ Px) + float(self.q @ self.x)
        return R

    def _prim_inf(self, dy, eps):        # src/auxil.c:364-427
        d = dy.copy()
        up_inf, lo_inf = self.u > OSQP_INFTY * MIN_SCALING, self.l < -OSQP_INFTY * MIN_SCALING
        d[up_inf & lo_inf] = 0.0
        only_up = up_inf & ~lo_inf
        d[only_up] = np.minimum(d[only_up], 0.0)
        only_lo = lo_inf & ~up_inf
        d[only_lo] = np.maximum(d[only_lo], 0.0)
        nrm = _inf(d)
        if not nrm > eps:
            return False
        lhs = float(self.u @ np.maximum(d, 0.0) + self.l @ np.minimum(d, 0.0))
        if not lhs < -eps * nrm:
            return False
        return _inf(self.A.T @ d) < eps * nrm

    def _dual_inf(self, dx, eps):        # src/auxil.c:429-515
        nrm = _inf(dx)
        if not nrm > eps:
            return False
        if not float(self.q @ dx) < -eps * nrm:
            return False
        if not _inf(self.P @ dx) < eps * nrm:
            return False
        Adx = self.A @ dx
        thr = eps * nrm
        bad = ((self.u < OSQP_INFTY * MIN_SCALING) & (Adx > thr)) | ((self.l > -OSQP_INFTY * MIN_SCALING) & (Adx < -thr))
        return not bool(np.any(bad))

    def _check(self, R, dx, dy, approximate):   # src/auxil.c:684-789
        s = self.s
        if R["pri"] > OSQP_INFTY or R["dua"] > OSQP_INFTY:
            return NON_CVX
        f = 10.0 if approximate else 1.0
        ea, er, epi, edi = f * s.eps_abs, f * s.eps_rel, f * s.eps_prim_inf, f * s.eps_dual_inf
        prim_ok = self.m == 0 or R["pri"] < ea + er * max(R["nz"], R["nAx"])
        prim_inf = (not prim_ok) and self._prim_inf(dy, epi)
        dual_ok = R["dua"] < ea + er * max(R["nq"], R["nAty"], R["nPx"])
        dual_inf = (not dual_ok) and self._dual_inf(dx, edi)
        if prim_ok and dual_ok:
            return SOLVED_INACC if approximate else SOLVED
        if prim_inf:
            return PRIM_INF_INACC if approximate else PRIM_INF
        if dual_inf:
            return DUAL_INF_INACC if approximate else DUAL_INF
        return None

    @staticmethod
    def _rho_estimate(R, rho):           # src/auxil.c:13-55
        pr = R["pri"] / (max(R["nz"], R["nAx"]) + 1e-10)
        dr = R["dua"] / (max(R["nq"], R["nAty"], R["nPx"]) + 1e-10)
        return min(max(rho * np.sqrt(pr / (dr + 1e-10)), RHO_MIN), RHO_MAX)

    def solve(self):
        """osqp_solve (src/osqp.c:288-641).  Returns a dict like OSQPInfo + solution."""
        s, n, m = self.s, self.n, self.m
        if not s.warm_start:
            self.x[:], self.z[:], self.y[:] = 0.0, 0.0, 0.0
        interval = s.adaptive_rho_interval
        if s.adaptive_rho and not interval:   # non-PROFILING rule, src/osqp.c:248-262
            interval = 4 * s.check_termination if s.check_termination else 100
        status, R, dx, dy = None, None, np.zeros(n), np.zeros(m)
        it = 0
        for it in range(1, s.max_iter + 1):
            xp, zp = self.x.copy(), self.z.copy()
            rinv = 1.0 / self.rho_vec
            rhs = np.concatenate([s.sigma * xp - self.q, zp - rinv * self.y])         # compute_rhs
            xz = self.kkt.solve(rhs)                                                  # update_xz_tilde
            self.x = s.alpha * xz[:n] + (1.0 - s.alpha) * xp                           # update_x
            dx = self.x - xp
            zh = s.alpha * xz[n:] + (1.0 - s.alpha) * zp
            self.z = np.minimum(np.maximum(zh + rinv * self.y, self.l), self.u)       # update_z + project
            dy = self.rho_vec * (zh - self.z)                                         # update_y
            self.y = self.y + dy
            can_check = bool(s.check_termination) and it % s.check_termination == 0
            R = None
            if can_check:
                R = self._residuals()
                status = self._check(R, dx, dy, False)
                if status is not None:
                    break
            if s.adaptive_rho and interval and it % interval == 0:
                if R is None:
                    R = self._residuals()
                rn = self._rho_estimate(R, self.rho)
                if rn > self.rho * s.adaptive_rho_tolerance or rn < self.rho / s.adaptive_rho_tolerance:
                    self.update_rho(rn)
                    self.rho_updates += 1
        if status is None:
            if R is None:
                R = self._residuals()
                status = self._check(R, dx, dy, False)
            if status is None:
                status = self._check(R, dx, dy, True)
                if status is None:
                    status = MAX_ITER
        no_sol = status in (PRIM_INF, PRIM_INF_INACC, DUAL_INF, DUAL_INF_INACC, NON_CVX)
        obj = R["obj"]
        if status in (PRIM_INF, PRIM_INF_INACC):
            obj = OSQP_INFTY
        if status in (DUAL_INF, DUAL_INF_INACC):
            obj = -OSQP_INFTY
        if status == NON_CVX:
            obj = OSQP_NAN
        out = dict(status=status, iter=it, obj=obj, pri_res=R["pri"], dua_res=R["dua"],
                   rho_updates=self.rho_updates, rho_estimate=self._rho_estimate(R, self.rho),
                   x=np.full(n, OSQP_NAN) if no_sol else self.x.copy(),
                   y=np.full(m, OSQP_NAN) if no_sol else self.y.copy())
        if no_sol:   # store_solution cold-starts the iterates (src/auxil.c:539-563)
            self.x[:], self.z[:], self.y[:] = 0.0, 0.0, 0.0
        return out


def mpc_solver(model, N, q, l, u, settings: Settings, path="recursive"):
    """Solver for one MPC instance of rldl_b200.mpc.MPCModel (q,l,u: 1-D, sized >= current n,m)."""
    blocks = model.dense_blocks()
    nx, nu, ny, nt = model.nx, model.nu, model.ny, model.nt
    Pu, A = model.assemble(N)
    Pu = Pu.toarray()
    P = Pu + Pu.T - np.diag(np.diag(Pu))
    A = A.toarray()
    n, m = P.shape[0], A.shape[0]
    kkt = StageKKT(blocks, nx, nu, ny, nt, N, settings.sigma) if path == "recursive" else DenseKKT(P, A, settings.sigma)
    return Solver(P, A, q[:n], l[:m], u[:m], kkt, settings)


class OracleBatch:
    """Batch facade over per-instance Solvers with the same surface as rldl_b200.Batch /
    oracle.ref_py.RefBatch (solve, get, dims, update_bounds, warm_start, update_horizon...)."""

    def __init__(self, model, N, Nmax, q, l, u, settings: dict, path="recursive", rho=None):
        keys = vars(Settings()).keys()
        self.model, self.N, self.Nmax, self.path = model, N, Nmax, path
        self.q, self.l, self.u = np.array(q, float), np.array(l, float), np.array(u, float)
        self.batch, self.nmax, self.mmax = q.shape[0], q.shape[1], l.shape[1]
        self.sv = []
        for k in range(self.batch):
            s = Settings(**{kk: vv for kk, vv in settings.items() if kk in keys})
            if rho is not None:
                s.rho = float(rho[k])
            self.sv.append(mpc_solver(model, N, self.q[k], self.l[k], self.u[k], s, path))
        self.last = None

    def dims(self):
        return self.sv[0].n, self.sv[0].m

    def solve(self):
        self.last = [s.solve() for s in self.sv]
        return 0

    def get(self):
        B, n, m = self.batch, *self.dims()
        out = dict(x=np.zeros((B, self.nmax)), y=np.zeros((B, self.mmax)))
        for key, dt in (("iter", np.int64), ("status", np.int64), ("obj", float), ("pri_res", float),
                        ("dua_res", float), ("rho_updates", np.int64), ("rho_estimate", float)):
            out[key] = np.array([r[key] for r in self.last], dt)
        for k, r in enumerate(self.last):
            out["x"][k, :n], out["y"][k, :m] = r["x"], r["y"]
        return out

    def update_bounds(self, l, u):
        self.l, self.u = np.array(l, float), np.array(u, float)
        for k, s in enumerate(self.sv):
            s.update_bounds(self.l[k], self.u[k])
        return 0

    def partial_update_bounds(self, start, stop, l, u):
        """osqp_partial_update_bounds (src/recursive_ldl.c:119): rows [start, stop) only."""
        ln, un = self.l.copy(), self.u.copy()
        ln[:, start:stop], un[:, start:stop] = l, u
        return self.update_bounds(ln, un)

    def update_lin_cost(self, q):
        self.q = np.array(q, float)
        for k, s in enumerate(self.sv):
            s.update_lin_cost(self.q[k])
        return 0

    def warm_start(self, x=None, y=None):
        for k, s in enumerate(self.sv):
            s.warm_start(None if x is None else x[k], None if y is None else y[k])
        return 0

    def update_rho(self, rho):
        for k, s in enumerate(self.sv):
            s.update_rho(float(rho[k]))
        return 0

    def update_horizon(self, N):
        """osqp_update_recursive (src/recursive_ldl.c:1973): new P/A sizes, tail re-factorisation."""
        Pu, A = self.model.assemble(N)
        Pu = Pu.toarray()
        P, A = Pu + Pu.T - np.diag(np.diag(Pu)), A.toarray()
        n, m = P.shape[0], A.shape[0]
        for k, s in enumerate(self.sv):
            xo, zo, yo = s.x, s.z, s.y
            s.P, s.A, s.n, s.m = P, A, n, m
            s.q, s.l, s.u = self.q[k, :n].copy(), self.l[k, :m].copy(), self.u[k, :m].copy()
            s.rho_vec = set_rho_vec(s.l, s.u, s.rho)
            if isinstance(s.kkt, StageKKT):
                s.kkt.set_horizon(N, s.rho_vec)
            else:
                s.kkt = DenseKKT(P, A, s.s.sigma).factor(s.rho_vec)
            s.x, s.z, s.y = np.zeros(n), np.zeros(m), np.zeros(m)
            s.x[:min(n, xo.size)] = xo[:n]; s.z[:min(m, zo.size)] = zo[:m]; s.y[:min(m, yo.size)] = yo[:m]
        self.N = N
        return 0

    def kkt_solve(self, k, rhs):
        return self.sv[k].kkt.solve(np.asarray(rhs, float))

    def free(self):
        pass
