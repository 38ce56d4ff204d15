"""ctypes binding of oracle/_ref/libref_driver.so — TEST INFRASTRUCTURE ONLY.

Runs the UNMODIFIED reference (compiled by oracle/Makefile from /root/reference)
over a batch of MPC instances. Imported only by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs; never by the product path.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_ref", "libref_driver.so")

c_int = C.c_longlong
c_float = C.c_double


class OSQPSettings(C.Structure):
    """Layout of reference include/types.h:139-176 for the default desktop build
    (EMBEDDED undefined, PROFILING defined, DLONG, double)."""
    _fields_ = [
        ("rho", c_float), ("sigma", c_float), ("scaling", c_int),
        ("adaptive_rho", c_int), ("adaptive_rho_interval", c_int),
        ("adaptive_rho_tolerance", c_float), ("adaptive_rho_fraction", c_float),
        ("max_iter", c_int), ("eps_abs", c_float), ("eps_rel", c_float),
        ("eps_prim_inf", c_float), ("eps_dual_inf", c_float), ("alpha", c_float),
        ("linsys_solver", C.c_int),
        ("delta", c_float), ("polish", c_int), ("polish_refine_iter", c_int), ("verbose", c_int),
        ("scaled_termination", c_int), ("check_termination", c_int), ("warm_start", c_int),
        ("time_limit", c_float),
    ]


def available() -> bool:
    return os.path.exists(_LIB)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_LIB)
        L.refb_create.restype = C.c_void_p
        L.refb_create.argtypes = [c_int] * 6 + [C.c_void_p] * 5 + [C.POINTER(OSQPSettings), c_int] + \
            [C.c_void_p] * 4 + [C.c_int, C.c_int]
        for name, args in [
            ("refb_solve", [C.c_void_p, C.c_int]),
            ("refb_get", [C.c_void_p] * 12),
            ("refb_update_bounds", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
            ("refb_partial_update_bounds", [C.c_void_p, c_int, c_int, C.c_void_p, C.c_void_p, C.c_int]),
            ("refb_update_lin_cost", [C.c_void_p, C.c_void_p, C.c_int]),
            ("refb_warm_start", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
            ("refb_update_horizon", [C.c_void_p, c_int, C.c_int]),
            ("refb_update_rho", [C.c_void_p, C.c_void_p, C.c_int]),
            ("refb_kkt_solve", [C.c_void_p, c_int, C.c_void_p]),
            ("refb_dims", [C.c_void_p, C.c_void_p, C.c_void_p]),
            ("refb_get_scaling", [C.c_void_p] * 4),
            ("refb_get_polish", [C.c_void_p, C.c_void_p]),
            ("refb_update_PA", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
            ("refb_get_PA", [C.c_void_p] * 5),
            ("refb_free", [C.c_void_p]),
            ("refb_default_settings", [C.POINTER(OSQPSettings)]),
        ]:
            getattr(L, name).argtypes = args
        assert L.refb_sizeof_settings() == C.sizeof(OSQPSettings)
        _lib = L
    return _lib


def default_settings(**kw) -> OSQPSettings:
    s = OSQPSettings()
    lib().refb_default_settings(C.byref(s))
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class RefBatch:
    """A batch of reference workspaces. path: 'recursive' | 'stock'."""

    def __init__(self, model, N, Nmax, q, l, u, settings: OSQPSettings, path="recursive",
                 rho=None, nthreads=None):
        L = lib()
        self.nthreads = nthreads or L.refb_max_threads()
        self.model, self.N, self.Nmax = model, N, Nmax
        mats = model.stage_matrices()
        self._keep = mats
        mm = np.array([t[0] for t in mats], np.int64)
        nn = np.array([t[1] for t in mats], np.int64)
        PP = (C.c_void_p * 7)(*[t[2].ctypes.data for t in mats])
        II = (C.c_void_p * 7)(*[t[3].ctypes.data for t in mats])
        XX = (C.c_void_p * 7)(*[t[4].ctypes.data for t in mats])
        self.batch = q.shape[0]
        self.nmax, self.mmax = q.shape[1], l.shape[1]
        q = np.ascontiguousarray(q, np.float64); l = np.ascontiguousarray(l, np.float64)
        u = np.ascontiguousarray(u, np.float64)
        rho = None if rho is None else np.ascontiguousarray(rho, np.float64)
        self.h = L.refb_create(model.nx, model.nu, model.ny, model.nt, N, Nmax, _p(mm), _p(nn),
                               C.cast(PP, C.c_void_p), C.cast(II, C.c_void_p), C.cast(XX, C.c_void_p),
                               C.byref(settings), self.batch, _p(q), _p(l), _p(u), _p(rho),
                               0 if path == "recursive" else 1, self.nthreads)

    def dims(self):
        n, m = c_int(), c_int()
        lib().refb_dims(self.h, C.byref(n), C.byref(m))
        return n.value, m.value

    def solve(self):
        return lib().refb_solve(self.h, self.nthreads)

    def get(self):
        B = self.batch
        out = dict(x=np.zeros((B, self.nmax)), y=np.zeros((B, self.mmax)),
                   iter=np.zeros(B, np.int64), status=np.zeros(B, np.int64), obj=np.zeros(B),
                   pri_res=np.zeros(B), dua_res=np.zeros(B), rho_updates=np.zeros(B, np.int64),
                   rho_estimate=np.zeros(B), setup_time=np.zeros(B), solve_time=np.zeros(B))
        lib().refb_get(self.h, *[_p(out[k]) for k in
                                 ("x", "y", "iter", "status", "obj", "pri_res", "dua_res",
                                  "rho_updates", "rho_estimate", "setup_time", "solve_time")])
        return out

    def update_bounds(self, l, u):
        return lib().refb_update_bounds(self.h, _p(np.ascontiguousarray(l)), _p(np.ascontiguousarray(u)), self.nthreads)

    def partial_update_bounds(self, start, stop, l, u):
        return lib().refb_partial_update_bounds(self.h, start, stop, _p(np.ascontiguousarray(l)),
                                                _p(np.ascontiguousarray(u)), self.nthreads)

    def update_lin_cost(self, q):
        return lib().refb_update_lin_cost(self.h, _p(np.ascontiguousarray(q)), self.nthreads)

    def warm_start(self, x=None, y=None):
        x = None if x is None else np.ascontiguousarray(x)
        y = None if y is None else np.ascontiguousarray(y)
        return lib().refb_warm_start(self.h, _p(x), _p(y), self.nthreads)

    def update_horizon(self, N):
        rc = lib().refb_update_horizon(self.h, N, self.nthreads)
        self.N = N
        return rc

    def update_rho(self, rho):
        return lib().refb_update_rho(self.h, _p(np.ascontiguousarray(rho, np.float64)), self.nthreads)

    def get_scaling(self):
        """(D [batch, nmax], E [batch, mmax], c [batch]) of the reference's equilibration, None if scaling is off."""
        D = np.zeros((self.batch, self.nmax)); E = np.zeros((self.batch, self.mmax)); c = np.zeros(self.batch)
        if lib().refb_get_scaling(self.h, _p(D), _p(E), _p(c)):
            return None
        return D, E, c

    def get_polish(self):
        sp = np.zeros(self.batch, np.int64)
        lib().refb_get_polish(self.h, _p(sp))
        return sp

    def get_PA(self):
        """Values of the reference's global P (upper triangle) and A of instance 0, CSC order (scaled if scaling is on)."""
        a, b = c_int(), c_int()
        lib().refb_get_PA(self.h, C.byref(a), C.byref(b), None, None)
        Px, Ax = np.zeros(a.value), np.zeros(b.value)
        lib().refb_get_PA(self.h, C.byref(a), C.byref(b), _p(Px), _p(Ax))
        return Px, Ax

    def update_PA(self, Px, Ax):
        return lib().refb_update_PA(self.h, _p(np.ascontiguousarray(Px, np.float64)),
                                    _p(np.ascontiguousarray(Ax, np.float64)), self.nthreads)

    def kkt_solve(self, k, rhs):
        rhs = np.ascontiguousarray(rhs, np.float64).copy()
        lib().refb_kkt_solve(self.h, k, _p(rhs))
        return rhs

    def free(self):
        if self.h:
            lib().refb_free(self.h)
            self.h = None


# ---------------------------------------------------------------------------
# Process isolation.  The reference's recursive path corrupts the heap on some
# shapes (L buffers sized from nnz(P), one write past L->p when
# N == Nmax), which can abort the *calling* process much later.  Tests and the
# bench therefore run every reference job in a throw-away child process.
# ---------------------------------------------------------------------------
def run_ops(spec):
    """Execute a scripted sequence on one RefBatch (in THIS process).

    spec = dict(model=MPCModel, N=, Nmax=, q=, l=, u=, settings=dict, path=, rho=None,
                nthreads=None, ops=[(name, *args), ...])
    ops: ("solve",) ("get",) ("update_bounds", l, u) ("partial_update_bounds", s, e, l, u)
         ("update_lin_cost", q) ("warm_start", x, y) ("update_horizon", N) ("update_rho", rho)
         ("kkt_solve", k, rhs) ("dims",)
    Returns a list with one entry per op (return code, dict or array) plus timing of
    the constructor as result[-1] = dict(setup_wall=, solve_wall=[...]).
    """
    import time
    t0 = time.perf_counter()
    rb = RefBatch(spec["model"], spec["N"], spec["Nmax"], spec["q"], spec["l"], spec["u"],
                  default_settings(**spec["settings"]), path=spec.get("path", "recursive"),
                  rho=spec.get("rho"), nthreads=spec.get("nthreads"))
    setup_wall = time.perf_counter() - t0
    res, solve_wall = [], []
    for op in spec["ops"]:
        name, args = op[0], op[1:]
        if name == "receding":       # ("receding", x0, T, shrink_every, feedback)
            from oracle import scenarios
            res.append(scenarios.receding_horizon(rb, spec["model"], spec["N"], spec["Nmax"], *args))
        elif name == "solve":
            t0 = time.perf_counter()
            res.append(rb.solve())
            solve_wall.append(time.perf_counter() - t0)
        else:
            res.append(getattr(rb, name)(*args))
    res.append(dict(setup_wall=setup_wall, solve_wall=solve_wall, nthreads=rb.nthreads))
    return res  # the batch is leaked on purpose: the process is about to exit


def run_isolated(spec, timeout=1800):
    """run_ops(spec) in a child interpreter (pickle in / pickle out through a temp dir)."""
    import pickle
    import subprocess
    import sys
    import tempfile
    root = os.path.dirname(_HERE)
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.pkl"), os.path.join(td, "out.pkl")
        with open(fin, "wb") as f:
            pickle.dump(spec, f, protocol=4)
        env = dict(os.environ)
        env["PYTHONPATH"] = os.pathsep.join([root, os.path.join(root, "osqp-recursive-ldl_b200"),
                                             env.get("PYTHONPATH", "")])
        cp = subprocess.run([sys.executable, "-m", "oracle.ref_worker", fin, fout], env=env, cwd=root,
                            timeout=timeout, capture_output=True, text=True)
        if not os.path.exists(fout):
            raise RuntimeError(f"reference worker failed (rc={cp.returncode}): {cp.stderr[-2000:]}")
        with open(fout, "rb") as f:
            return pickle.load(f)
