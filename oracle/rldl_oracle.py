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
  ruiz_scale ............ scale_data, src/scaling.c:44-156; with settings.scaling > 0 the Solver holds the
                          SCALED data and iterates, as the reference does, unscales residuals / tolerances /
                          certificates per src/auxil.c:251-515 and the solution per scaling.c:177-192
                          (pinned against the reference's stock path with scaling = 10: tests/test_oracle_cpu.py)
  Solver.polish ......... polish, src/polish.c:212-350 (reduced KKT of the active constraints, iterative
                          refinement :120-168, acceptance test :300-330)
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
        self.scaling, self.polish, self.delta, self.polish_refine_iter = 10, 0, 1e-6, 3
        for k, v in kw.items():
            setattr(self, k, v)


class Solver:
    """One QP instance: min 1/2 x'Px + q'x  s.t. l <= Ax <= u, with a pluggable KKT object
    (StageKKT for the recursive path, DenseKKT for the generic path)."""

    def __init__(self, P, A, q, l, u, kkt, settings: Settings):
        self.P, self.A = np.asarray(P, float), np.asarray(A, float)   # P full symmetric
        self.q, self.l, self.u = np.array(q, float), np.array(l, float), np.array(u, float)
        self.s, self.kkt = settings, kkt
        self.n, self.m = self.P.shape[0], self.A.shape[0]
        self.P0, self.A0 = self.P, self.A                               # the unscaled problem (polish works on it)
        self.D, self.E, self.c = np.ones(self.n), np.ones(self.m), 1.0
        if settings.scaling:
            # scale_data (src/scaling.c:44): from here on P, A, q, l, u, x, z, y are the SCALED quantities, as
            # in the reference's workspace; the reference's stock path factors the scaled KKT (src/osqp.c:167-216)
            self.D, self.E, self.c, self.P, self.A, self.q = ruiz_scale(self.P, self.A, self.q, settings.scaling)
            self.l, self.u = self.E * self.l, self.E * self.u
            self.kkt = DenseKKT(self.P, self.A, settings.sigma)
        self.rho = min(max(settings.rho, RHO_MIN), RHO_MAX)
        self.rho_vec = set_rho_vec(self.l, self.u, self.rho)
        self.kkt.factor(self.rho_vec)
        self.x, self.z, self.y = np.zeros(self.n), np.zeros(self.m), np.zeros(self.m)
        self.rho_updates = 0

    # osqp_warm_start (src/osqp.c:929-994)
    def warm_start(self, x=None, y=None):
        self.s.warm_start = 1
        if x is not None:
            self.x = np.array(x[: self.n], float) / self.D              # osqp.c:942-943
            self.z = self.A @ self.x
        if y is not None:
            self.y = np.array(y[: self.m], float) / self.E * self.c     # osqp.c:944-945

    # osqp_update_bounds (src/osqp.c:784) -> update_rho_vec (src/auxil.c:103-145)
    def update_bounds(self, l, u):
        self.l, self.u = self.E * np.array(l[: self.m], float), self.E * np.array(u[: self.m], float)  # osqp.c:815-817
        new = set_rho_vec(self.l, self.u, self.rho)
        if np.any(new != self.rho_vec):
            self.rho_vec = new
            self.kkt.factor(self.rho_vec)

    def update_lin_cost(self, q):
        self.q = self.c * self.D * np.array(q[: self.n], float)         # osqp.c:769-771

    # osqp_update_rho (src/osqp.c:1268-1319)
    def update_rho(self, rho_new):
        self.rho = min(max(rho_new, RHO_MIN), RHO_MAX)
        self.rho_vec = set_rho_vec(self.l, self.u, self.rho)
        self.kkt.factor(self.rho_vec)

    def _residuals(self, x=None, z=None, y=None):
        # compute_pri_res / compute_dua_res and their tolerances' normalisers, unscaled as auxil.c:251-345 does
        # (scaled_termination = 0); "s": the norms of the scaled vectors, which compute_rho_estimate reads
        x, z, y = (self.x if x is None else x), (self.z if z is None else z), (self.y if y is None else y)
        Ax, Px, Aty = self.A @ x, self.P @ x, self.A.T @ y
        Ei, Di, ci = 1.0 / self.E, 1.0 / self.D, 1.0 / self.c
        R = dict(pri=_inf(Ei * (Ax - z)), nz=_inf(Ei * z), nAx=_inf(Ei * Ax),
                 dua=ci * _inf(Di * (self.q + Px + Aty)), nq=ci * _inf(Di * self.q), nAty=ci * _inf(Di * Aty),
                 nPx=ci * _inf(Di * Px))
        R["s"] = dict(pri=_inf(Ax - z), nz=_inf(z), nAx=_inf(Ax), dua=_inf(self.q + Px + Aty),
                      nq=_inf(self.q), nAty=_inf(Aty), nPx=_inf(Px))
        R["obj"] = ci * (0.5 * float(x @ Px) + float(self.q @ x))
        return R

    def _polish(self, info):
        """polish, src/polish.c:212-350, on the (scaled) workspace data.  Updates info and the iterates in place
        when the polished point is accepted; returns status_polish (1 accepted, -1 rejected)."""
        s, n = self.s, self.n
        low = (self.z - self.l) < -self.y                                   # form_Ared :35-49
        upp = ~low & ((self.u - self.z) < self.y)
        il, iu = np.flatnonzero(low), np.flatnonzero(upp)
        Ared = np.vstack([self.A[il], self.A[iu]]) if len(il) + len(iu) else np.zeros((0, n))
        mr = Ared.shape[0]
        K = np.block([[self.P, Ared.T], [Ared, np.zeros((mr, mr))]])
        Kreg = K + np.diag(np.concatenate([np.full(n, s.delta), np.full(mr, -s.delta)]))   # kkt.c form_KKT, polish
        b = np.concatenate([-self.q, self.l[il], self.u[iu]])                # form_rhs_red :103-118
        lu = sla.lu_factor(Kreg)
        zz = sla.lu_solve(lu, b)
        for _ in range(int(s.polish_refine_iter)):                          # iterative_refinement :120-168
            zz = zz + sla.lu_solve(lu, b - K @ zz)
        x = zz[:n]
        zp = self.A @ x
        yp = np.zeros(self.m)                                               # get_ypol_from_yred :176-201
        yp[il], yp[iu] = zz[n:n + len(il)], zz[n + len(il):]
        t = zp + yp                                                         # project_normalcone, proj.c:16-29
        zp = np.minimum(np.maximum(t, self.l), self.u)
        yp = t - zp
        R = self._residuals(x, zp, yp)
        if self.m == 0:
            R["pri"] = 0.0
        ok = (R["pri"] < info["pri_res"] and R["dua"] < info["dua_res"]) or \
             (R["pri"] < info["pri_res"] and info["dua_res"] < 1e-10) or \
             (R["dua"] < info["dua_res"] and info["pri_res"] < 1e-10)       # :300-316
        if ok:
            info.update(obj=R["obj"], pri_res=R["pri"], dua_res=R["dua"])
            self.x, self.z, self.y = x, zp, yp
        return 1 if ok else -1

    def _prim_inf(self, dy, eps):        # src/auxil.c:364-427
        d = dy.copy()
        up_inf, lo_inf = self.u > OSQP_INFTY * MIN_SCALING, self.l < -OSQP_INFTY * MIN_SCALING
        d[up_inf & lo_inf] = 0.0
        only_up = up_inf & ~lo_inf
        d[only_up] = np.minimum(d[only_up], 0.0)
        only_lo = lo_inf & ~up_inf
        d[only_lo] = np.maximum(d[only_lo], 0.0)
        nrm = _inf(self.E * d)
        if not nrm > eps:
            return False
        lhs = float(self.u @ np.maximum(d, 0.0) + self.l @ np.minimum(d, 0.0))
        if not lhs < -eps * nrm:
            return False
        return _inf((self.A.T @ d) / self.D) < eps * nrm

    def _dual_inf(self, dx, eps):        # src/auxil.c:429-515
        nrm = _inf(self.D * dx)
        if not nrm > eps:
            return False
        if not float(self.q @ dx) < -self.c * eps * nrm:
            return False
        if not _inf((self.P @ dx) / self.D) < self.c * eps * nrm:
            return False
        Adx = (self.A @ dx) / self.E
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
    def _rho_estimate(R, rho):           # src/auxil.c:13-55 (on the scaled vectors)
        R = R.get("s", R)
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
                   rho_updates=self.rho_updates, rho_estimate=self._rho_estimate(R, self.rho), status_polish=0)
        if s.polish and status == SOLVED:                                   # osqp.c:589-591
            out["status_polish"] = self._polish(out)
        out.update(x=np.full(n, OSQP_NAN) if no_sol else self.D * self.x,                 # unscale_solution,
                   y=np.full(m, OSQP_NAN) if no_sol else self.E * self.y / self.c)       # scaling.c:177-192
        if no_sol:   # store_solution cold-starts the iterates (src/auxil.c:539-563)
            self.x[:], self.z[:], self.y[:] = 0.0, 0.0, 0.0
        return out


def _limit_scaling(v):
    """src/scaling.c:7-14: norms below MIN_SCALING count as 1, above MAX_SCALING as MAX_SCALING."""
    v = np.where(v < MIN_SCALING, 1.0, v)
    return np.minimum(v, 1e4)


def ruiz_scale(P, A, q, iters):
    """scale_data, src/scaling.c:44-156.  P full symmetric.  Returns D, E, c and the scaled P, A, q."""
    P, A, q = np.array(P, float), np.array(A, float), np.array(q, float)
    n, m = P.shape[0], A.shape[0]
    D, E, c = np.ones(n), np.ones(m), 1.0
    for _ in range(int(iters)):
        dn = np.abs(P).max(axis=0) if n else np.zeros(0)
        if m:
            dn = np.maximum(dn, np.abs(A).max(axis=0))
            en = np.abs(A).max(axis=1)
        else:
            en = np.zeros(0)
        dt, et = 1.0 / np.sqrt(_limit_scaling(dn)), 1.0 / np.sqrt(_limit_scaling(en))
        P = dt[:, None] * P * dt[None, :]
        A = et[:, None] * A * dt[None, :]
        q = dt * q
        D, E = D * dt, E * et
        ct = max(float(np.abs(P).max(axis=0).mean()), float(_limit_scaling(np.array(_inf(q)))))
        ct = 1.0 / float(_limit_scaling(np.array(ct)))
        P, q, c = P * ct, q * ct, c * ct
    return D, E, c, P, A, q


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
                        ("dua_res", float), ("rho_updates", np.int64), ("rho_estimate", float),
                        ("status_polish", np.int64)):
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
            assert not s.s.scaling, "horizon changes are restated for the unscaled recursive path only"
            s.P, s.A, s.n, s.m = P, A, n, m
            s.P0, s.A0, s.D, s.E = P, A, np.ones(n), np.ones(m)
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
