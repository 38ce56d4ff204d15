"""Synthetic linear-MPC QP instances in the reference's stage-block input format.

One definition shared by the oracle legs and the B200 path. The
stage matrices follow the layout `osqp_setup_recursive` expects
(reference include/recursive_ldl.h:17-50, src/recursive_ldl.c:1873-1971):

  variables    x = [u0 | (x1,u1) | ... | (x_{N-1},u_{N-1}) | xN]        n = N*(nx+nu)
  constraints  c_k = [C (x_k,u_k) ; A x_k + B u_k - x_{k+1}], k<N ; cN  m = N*(nx+ny)+nt

with ny = nx+nu, C = I (box on state and input), nt = nx, AN = I.
Cost blocks are stored as FULL upper triangles with explicit zeros
so every diagonal entry exists (sigma is only added to stored diagonals,
reference src/cs_addon.c:65).
"""
from __future__ import annotations

import dataclasses
import numpy as np
import scipy.sparse as sp

INF = 1e30  # OSQP_INFTY, reference include/constants.h


def _csc_full(M: np.ndarray, upper: bool = False):
    """Dense -> CSC keeping explicit zeros (all entries, or the upper triangle)."""
    m, n = M.shape
    p = [0]
    idx, val = [], []
    for j in range(n):
        rows = range(0, min(j + 1, m)) if upper else range(m)
        for i in rows:
            idx.append(i)
            val.append(M[i, j])
        p.append(len(idx))
    return (m, n, np.asarray(p, np.int64), np.asarray(idx, np.int64), np.asarray(val, np.float64))


def _csc_sparse(M: np.ndarray):
    S = sp.csc_matrix(M)
    S.sort_indices()
    return (M.shape[0], M.shape[1], S.indptr.astype(np.int64), S.indices.astype(np.int64),
            S.data.astype(np.float64))


@dataclasses.dataclass
class MPCModel:
    nx: int
    nu: int
    A: np.ndarray
    B: np.ndarray
    Q: np.ndarray
    R: np.ndarray
    xmin: np.ndarray
    xmax: np.ndarray
    umin: np.ndarray
    umax: np.ndarray
    xr: np.ndarray
    x0_scale: np.ndarray | float = 1.0

    @property
    def ny(self):
        return self.nx + self.nu

    @property
    def nt(self):
        return self.nx

    def stage_matrices(self):
        """(Q0,Qi,QN,A0,Ai,Aij,AN) as (m,n,p,i,x) CSC tuples in the reference's layout."""
        nx, nu = self.nx, self.nu
        a, b = nx + nu, nx + self.ny
        Qi = np.zeros((a, a)); Qi[:nx, :nx] = self.Q; Qi[nx:, nx:] = self.R
        A0 = np.zeros((b, nu)); A0[nx:nx + nu, :] = np.eye(nu); A0[self.ny:, :] = self.B
        Ai = np.zeros((b, a)); Ai[:a, :] = np.eye(a); Ai[self.ny:, :nx] = self.A; Ai[self.ny:, nx:] = self.B
        Aij = np.zeros((b, a)); Aij[self.ny:, :nx] = -np.eye(nx)
        AN = np.eye(nx)
        return [_csc_full(self.R, True), _csc_full(Qi, True), _csc_full(self.Q, True),
                _csc_sparse(A0), _csc_sparse(Ai), _csc_sparse(Aij), _csc_sparse(AN)]

    def dense_blocks(self):
        nx, nu = self.nx, self.nu
        a, b = nx + nu, nx + self.ny
        Qi = np.zeros((a, a)); Qi[:nx, :nx] = self.Q; Qi[nx:, nx:] = self.R
        A0 = np.zeros((b, nu)); A0[nx:nx + nu, :] = np.eye(nu); A0[self.ny:, :] = self.B
        Ai = np.zeros((b, a)); Ai[:a, :] = np.eye(a); Ai[self.ny:, :nx] = self.A; Ai[self.ny:, nx:] = self.B
        Aij = np.zeros((b, a)); Aij[self.ny:, :nx] = -np.eye(nx)
        return dict(Q0=self.R.copy(), Qi=Qi, QN=self.Q.copy(), A0=A0, Ai=Ai, Aij=Aij, AN=np.eye(nx))

    def vectors(self, x0: np.ndarray, N: int, Nmax: int | None = None):
        """q,l,u for a batch of initial states x0 [batch,nx]; sized for Nmax and
        pre-filled stage-like up to Nmax."""
        Nmax = N if Nmax is None else Nmax
        x0 = np.atleast_2d(np.asarray(x0, np.float64))
        B_ = x0.shape[0]
        nx, nu, ny = self.nx, self.nu, self.ny
        a, b = nx + nu, nx + ny
        n, m = Nmax * a, Nmax * b + nx
        q = np.zeros((B_, n)); l = np.zeros((B_, m)); u = np.zeros((B_, m))
        qx = -self.Q @ self.xr
        for k in range(1, Nmax + 1):          # x_k lives at nu + (k-1)*a
            off = nu + (k - 1) * a
            q[:, off:off + nx] = qx
        for k in range(Nmax):
            r = k * b
            if k == 0:
                l[:, r:r + nx] = -INF; u[:, r:r + nx] = INF
            else:
                l[:, r:r + nx] = self.xmin; u[:, r:r + nx] = self.xmax
            l[:, r + nx:r + ny] = self.umin; u[:, r + nx:r + ny] = self.umax
            # dynamics rows: 0, except stage 0: -A x0
        dyn0 = -(x0 @ self.A.T)
        l[:, ny:ny + nx] = dyn0; u[:, ny:ny + nx] = dyn0
        # terminal rows for the *current* N and for Nmax
        for NN in {N, Nmax}:
            r = NN * b
            l[:, r:r + nx] = self.xmin; u[:, r:r + nx] = self.xmax
        return q, l, u

    def assemble(self, N: int):
        """Global (P upper-tri, A) as scipy CSC — mirrors reference setup_AP_matrices
        (src/recursive_ldl.c:1873) for test-side cross checks."""
        d = self.dense_blocks()
        nx, nu = self.nx, self.nu
        a, b = nx + nu, nx + self.ny
        n, m = N * a, N * b + nx
        P = np.zeros((n, n)); A = np.zeros((m, n))
        P[:nu, :nu] = d["Q0"]; A[:b, :nu] = d["A0"]
        for k in range(1, N):
            c = nu + (k - 1) * a
            P[c:c + a, c:c + a] = d["Qi"]
            A[(k - 1) * b:k * b, c:c + a] += d["Aij"]
            A[k * b:(k + 1) * b, c:c + a] += d["Ai"]
        c = nu + (N - 1) * a
        P[c:c + nx, c:c + nx] = d["QN"]
        A[(N - 1) * b:N * b, c:c + nx] += d["Aij"][:, :nx]
        A[N * b:N * b + nx, c:c + nx] += d["AN"]
        return sp.csc_matrix(np.triu(P)), sp.csc_matrix(A)


def random_model(nx: int, nu: int, seed: int = 1234) -> MPCModel:
    """C1/C2/C4 model family: A = 0.9 I + 0.1 U(-.5,.5), B = U(-.5,.5)."""
    rng = np.random.default_rng(seed)
    A = 0.9 * np.eye(nx) + 0.1 * rng.uniform(-0.5, 0.5, (nx, nx))
    B = rng.uniform(-0.5, 0.5, (nx, nu))
    def cost(k, d):
        i, j = np.indices((k, k))
        M = 0.01 / (1.0 + np.abs(i - j))
        M[np.arange(k), np.arange(k)] = d
        return M
    return MPCModel(nx, nu, A, B, cost(nx, 1.0), cost(nu, 0.1),
                    -5.0 * np.ones(nx), 5.0 * np.ones(nx), -0.5 * np.ones(nu), 0.5 * np.ones(nu),
                    np.zeros(nx))


def quadcopter_model() -> MPCModel:
    """The 12-state / 4-input quadcopter of the reference docs (docs/examples/mpc.rst:31-76)."""
    Ad = np.array([
        [1., 0., 0., 0., 0., 0., 0.1, 0., 0., 0., 0., 0.],
        [0., 1., 0., 0., 0., 0., 0., 0.1, 0., 0., 0., 0.],
        [0., 0., 1., 0., 0., 0., 0., 0., 0.1, 0., 0., 0.],
        [0.0488, 0., 0., 1., 0., 0., 0.0016, 0., 0., 0.0992, 0., 0.],
        [0., -0.0488, 0., 0., 1., 0., 0., -0.0016, 0., 0., 0.0992, 0.],
        [0., 0., 0., 0., 0., 1., 0., 0., 0., 0., 0., 0.0992],
        [0., 0., 0., 0., 0., 0., 1., 0., 0., 0., 0., 0.],
        [0., 0., 0., 0., 0., 0., 0., 1., 0., 0., 0., 0.],
        [0., 0., 0., 0., 0., 0., 0., 0., 1., 0., 0., 0.],
        [0.9734, 0., 0., 0., 0., 0., 0.0488, 0., 0., 0.9846, 0., 0.],
        [0., -0.9734, 0., 0., 0., 0., 0., -0.0488, 0., 0., 0.9846, 0.],
        [0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0.9846]])
    Bd = np.array([
        [0., -0.0726, 0., 0.0726],
        [-0.0726, 0., 0.0726, 0.],
        [-0.0152, 0.0152, -0.0152, 0.0152],
        [-0., -0.0006, -0., 0.0006],
        [0.0006, 0., -0.0006, 0.0000],
        [0.0106, 0.0106, 0.0106, 0.0106],
        [0, -1.4512, 0., 1.4512],
        [-1.4512, 0., 1.4512, 0.],
        [-0.3049, 0.3049, -0.3049, 0.3049],
        [-0., -0.0236, 0., 0.0236],
        [0.0236, 0., -0.0236, 0.],
        [0.2107, 0.2107, 0.2107, 0.2107]])
    u0 = 10.5916
    umin = np.array([9.6] * 4) - u0
    umax = np.array([13.] * 4) - u0
    xmin = np.array([-np.pi / 6, -np.pi / 6, -INF, -INF, -INF, -1.] + [-INF] * 6)
    xmax = np.array([np.pi / 6, np.pi / 6] + [INF] * 10)
    Q = np.diag([0., 0., 10., 10., 10., 10., 0., 0., 0., 5., 5., 5.])
    R = 0.1 * np.eye(4)
    xr = np.array([0., 0., 1.] + [0.] * 9)
    x0_scale = np.array([0.2, 0.2, 0.5, 0.5, 0.5, 0.5, 0.3, 0.3, 0.3, 0.3, 0.3, 0.3])
    return MPCModel(12, 4, Ad, Bd, Q, R, xmin, xmax, umin, umax, xr, x0_scale)


def sample_x0(model: MPCModel, batch: int, seed: int = 1234) -> np.ndarray:
    """Per-instance initial states. Random family: U(-1,1)^nx. Quadcopter: small
    perturbations around hover so every instance stays feasible."""
    rng = np.random.default_rng(seed)
    return rng.uniform(-1.0, 1.0, (batch, model.nx)) * model.x0_scale


# Named configurations of BASELINE.json (C1..C5)
CONFIGS = {
    "C1": dict(model="random", nx=4, nu=2, N=20, batch=1),
    "C2": dict(model="random", nx=4, nu=2, N=20, batch=1024),
    "C3": dict(model="quadcopter", nx=12, nu=4, N=50, batch=4096),
    "C4": dict(model="random", nx=8, nu=3, N=100, batch=2048),
    "C5": dict(model="quadcopter", nx=12, nu=4, N=50, batch=65536),
}


def make_model(cfg: dict) -> MPCModel:
    if cfg["model"] == "quadcopter":
        return quadcopter_model()
    return random_model(cfg["nx"], cfg["nu"])


@dataclasses.dataclass
class PolytopicModel(MPCModel):
    """General stage blocks the reference's entry point admits (include/recursive_ldl.h:17-50):
    ny polytopic path rows C (x_k,u_k) with several variables per row, nt terminal rows CN x_N,
    and a stage cost with state-input cross terms S.  Stage 0 keeps the input columns of C
    (x0 is data).  xmin/xmax/umin/umax of the base class are unused."""
    C: np.ndarray = None       # ny x (nx+nu)
    CN: np.ndarray = None      # nt x nx
    S: np.ndarray = None       # nx x nu cross cost
    cmin: np.ndarray = None
    cmax: np.ndarray = None
    tmin: np.ndarray = None
    tmax: np.ndarray = None

    @property
    def ny(self):
        return self.C.shape[0]

    @property
    def nt(self):
        return self.CN.shape[0]

    def dense_blocks(self):
        nx, nu, ny = self.nx, self.nu, self.ny
        a, b = nx + nu, nx + ny
        Qi = np.zeros((a, a)); Qi[:nx, :nx] = self.Q; Qi[nx:, nx:] = self.R
        Qi[:nx, nx:] = self.S; Qi[nx:, :nx] = self.S.T
        A0 = np.zeros((b, nu)); A0[:ny, :] = self.C[:, nx:]; A0[ny:, :] = self.B
        Ai = np.zeros((b, a)); Ai[:ny, :] = self.C; Ai[ny:, :nx] = self.A; Ai[ny:, nx:] = self.B
        Aij = np.zeros((b, a)); Aij[ny:, :nx] = -np.eye(nx)
        return dict(Q0=self.R.copy(), Qi=Qi, QN=self.Q.copy(), A0=A0, Ai=Ai, Aij=Aij, AN=self.CN.copy())

    def stage_matrices(self):
        d = self.dense_blocks()
        return [_csc_full(d["Q0"], True), _csc_full(d["Qi"], True), _csc_full(d["QN"], True),
                _csc_sparse(d["A0"]), _csc_sparse(d["Ai"]), _csc_sparse(d["Aij"]), _csc_sparse(d["AN"])]

    def vectors(self, x0, N, Nmax=None):
        Nmax = N if Nmax is None else Nmax
        x0 = np.atleast_2d(np.asarray(x0, np.float64))
        B_ = x0.shape[0]
        nx, nu, ny, nt = self.nx, self.nu, self.ny, self.nt
        a, b = nx + nu, nx + ny
        n, m = Nmax * a, Nmax * b + nt
        q = np.zeros((B_, n)); l = np.zeros((B_, m)); u = np.zeros((B_, m))
        qx = -self.Q @ self.xr
        for k in range(1, Nmax + 1):
            off = nu + (k - 1) * a
            q[:, off:off + nx] = qx
        for k in range(Nmax):
            r = k * b
            l[:, r:r + ny] = self.cmin; u[:, r:r + ny] = self.cmax
        dyn0 = -(x0 @ self.A.T)
        l[:, ny:ny + nx] = dyn0; u[:, ny:ny + nx] = dyn0
        for NN in sorted({N, Nmax}, reverse=True):   # the current terminal rows last
            r = NN * b
            l[:, r:r + nt] = self.tmin; u[:, r:r + nt] = self.tmax
        return q, l, u

    def assemble(self, N):
        d = self.dense_blocks()
        nx, nu, nt = self.nx, self.nu, self.nt
        a, b = nx + nu, nx + self.ny
        n, m = N * a, N * b + nt
        P = np.zeros((n, n)); A = np.zeros((m, n))
        P[:nu, :nu] = d["Q0"]; A[:b, :nu] = d["A0"]
        for k in range(1, N):
            c = nu + (k - 1) * a
            P[c:c + a, c:c + a] = d["Qi"]
            A[(k - 1) * b:k * b, c:c + a] += d["Aij"]
            A[k * b:(k + 1) * b, c:c + a] += d["Ai"]
        c = nu + (N - 1) * a
        P[c:c + nx, c:c + nx] = d["QN"]
        A[(N - 1) * b:N * b, c:c + nx] += d["Aij"][:, :nx]
        A[N * b:N * b + nt, c:c + nx] += d["AN"]
        return sp.csc_matrix(np.triu(P)), sp.csc_matrix(A)


def polytopic_model(nx: int, nu: int, ny: int, nt: int, seed: int = 4321) -> PolytopicModel:
    """Random stable-ish dynamics, dense polytopic path/terminal rows, cross-term cost.  The LAST path
    row is loose (the reference types its terminal rows from the first nt rows of stage Nmax-1, SURVEY
    §8c F3, so those must be plain inequalities like the terminal rows for its recursive path to be right)."""
    base = random_model(nx, nu, seed)
    rng = np.random.default_rng(seed + 1)
    a = nx + nu
    Cm = rng.uniform(-1.0, 1.0, (ny, a))
    CN = rng.uniform(-1.0, 1.0, (nt, nx))
    S = 0.02 * rng.uniform(-1.0, 1.0, (nx, nu))
    cmin = -rng.uniform(1.0, 3.0, ny); cmax = rng.uniform(1.0, 3.0, ny)
    cmin[-1] = -INF; cmax[-1] = INF
    tmin = -rng.uniform(1.0, 2.0, nt); tmax = rng.uniform(1.0, 2.0, nt)
    return PolytopicModel(nx, nu, base.A, base.B, base.Q, base.R, base.xmin, base.xmax, base.umin, base.umax,
                          0.3 * np.ones(nx), 0.5, C=Cm, CN=CN, S=S, cmin=cmin, cmax=cmax, tmin=tmin, tmax=tmax)
