"""ctypes mirror of the C ABI in include/osqp_rldl_abi.h (librldl_b200.so).

Thin plumbing only: structs that match the reference's layout, and a `Batch`
convenience wrapper over the additive rldl_batch_* entry points.  There is no
Python or CPU compute path: if the CUDA library is missing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

from . import mpc  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
# RLDL_B200_LIB selects another build of the same library (the -DRLDL_DEBUG build with in-kernel asserts)
LIB_PATH = os.environ.get("RLDL_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "librldl_b200.so")

c_int = C.c_longlong
c_float = C.c_double


class csc(C.Structure):
    _fields_ = [("nzmax", c_int), ("m", c_int), ("n", c_int), ("p", C.POINTER(c_int)),
                ("i", C.POINTER(c_int)), ("x", C.POINTER(c_float)), ("nz", c_int)]


class OSQPSolution(C.Structure):
    _fields_ = [("x", C.POINTER(c_float)), ("y", C.POINTER(c_float))]


class OSQPInfo(C.Structure):
    _fields_ = [("iter", c_int), ("status", C.c_char * 32), ("status_val", c_int),
                ("status_polish", c_int), ("obj_val", c_float), ("pri_res", c_float),
                ("dua_res", c_float), ("setup_time", c_float), ("solve_time", c_float),
                ("update_time", c_float), ("polish_time", c_float), ("run_time", c_float),
                ("rho_updates", c_int), ("rho_estimate", c_float)]


class OSQPData(C.Structure):
    _fields_ = [("n", c_int), ("m", c_int), ("P", C.POINTER(csc)), ("A", C.POINTER(csc)),
                ("q", C.POINTER(c_float)), ("l", C.POINTER(c_float)), ("u", C.POINTER(c_float))]


class OSQPSettings(C.Structure):
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


class LinSysSolver(C.Structure):
    pass


LinSysSolver._fields_ = [
    ("type", C.c_int),
    ("solve", C.CFUNCTYPE(c_int, C.POINTER(LinSysSolver), C.POINTER(c_float))),
    ("free", C.CFUNCTYPE(None, C.POINTER(LinSysSolver))),
    ("update_matrices", C.CFUNCTYPE(c_int, C.POINTER(LinSysSolver), C.POINTER(csc), C.POINTER(csc))),
    ("update_rho_vec", C.CFUNCTYPE(c_int, C.POINTER(LinSysSolver), C.POINTER(c_float))),
    ("nthreads", c_int),
]

_pf = C.POINTER(c_float)


class OSQPWorkspace(C.Structure):
    _fields_ = [("data", C.POINTER(OSQPData)), ("linsys_solver", C.POINTER(LinSysSolver)),
                ("pol", C.c_void_p), ("rho_vec", _pf), ("rho_inv_vec", _pf),
                ("constr_type", C.POINTER(c_int)), ("x", _pf), ("y", _pf), ("z", _pf),
                ("xz_tilde", _pf), ("x_prev", _pf), ("z_prev", _pf), ("Ax", _pf), ("Px", _pf),
                ("Aty", _pf), ("delta_y", _pf), ("Atdelta_y", _pf), ("delta_x", _pf),
                ("Pdelta_x", _pf), ("Adelta_x", _pf), ("D_temp", _pf), ("D_temp_A", _pf),
                ("E_temp", _pf), ("settings", C.POINTER(OSQPSettings)), ("scaling", C.c_void_p),
                ("solution", C.POINTER(OSQPSolution)), ("info", C.POINTER(OSQPInfo)),
                ("timer", C.c_void_p), ("first_run", c_int), ("clear_update_time", c_int),
                ("rho_update_from_solve", c_int), ("summary_printed", c_int)]


class OSQPDataRLDL(C.Structure):
    _fields_ = [("data", OSQPData), ("Nmax", c_int), ("N", c_int), ("N0", c_int),
                ("N_recursive_temp", c_int), ("Nx", c_int), ("Nplus", c_int), ("nx", c_int),
                ("ny", c_int), ("nu", c_int), ("nt", c_int), ("nnz", c_int * 2),
                ("A0", C.POINTER(csc)), ("Ai", C.POINTER(csc)), ("Aij", C.POINTER(csc)),
                ("AN", C.POINTER(csc)), ("Q0", C.POINTER(csc)), ("Qi", C.POINTER(csc)),
                ("QN", C.POINTER(csc)), ("Amats", C.c_void_p), ("Qmats", C.c_void_p),
                ("X_even", C.c_void_p), ("Ytemp", C.c_void_p), ("Yhat", C.c_void_p),
                ("Uhat", C.c_void_p), ("Vhat", C.c_void_p), ("Lz", C.c_void_p),
                ("Dzinv", C.c_void_p), ("Pz", C.c_void_p), ("Ly", C.c_void_p),
                ("Dyinv", C.c_void_p), ("P0", C.c_void_p)]


# every symbol include/osqp_rldl_abi.h declares
ABI_SYMBOLS = """
osqp_version csc_matrix csc_spalloc csc_spfree copy_csc_mat prea_copy_csc_mat triplet_to_csc csc_to_triu
osqp_set_default_settings osqp_setup osqp_solve osqp_cleanup osqp_update_lin_cost osqp_update_bounds
osqp_update_lower_bound osqp_update_upper_bound osqp_warm_start osqp_warm_start_x osqp_warm_start_y
osqp_update_P osqp_update_A osqp_update_P_A osqp_update_rho osqp_update_max_iter osqp_update_eps_abs
osqp_update_eps_rel osqp_update_eps_prim_inf osqp_update_eps_dual_inf osqp_update_alpha
osqp_update_warm_start osqp_update_scaled_termination osqp_update_check_termination osqp_update_delta
osqp_update_polish osqp_update_polish_refine_iter osqp_update_verbose osqp_update_time_limit
osqp_setup_recursive osqp_solve_recursive osqp_update_recursive osqp_setup_combine_recursive
osqp_solve_combine_recursive osqp_update_Z_horizon osqp_update_X_horizon
osqp_allocate_combine_recursive osqp_partial_update_bounds get_L_dimensions cleanup_rldl
rldl_batch_setup rldl_batch_solve rldl_batch_solve_to_host rldl_batch_update_bounds rldl_batch_partial_update_bounds
rldl_batch_update_lin_cost rldl_batch_update_data_refactor rldl_batch_warm_start rldl_batch_warm_start_shift rldl_batch_update_horizon rldl_batch_update_rho
rldl_batch_refactor rldl_batch_update_model rldl_batch_failed_factorisations rldl_batch_kkt_solve rldl_batch_kkt_solve_static rldl_batch_max_static_doubles rldl_batch_get_solution rldl_batch_get_info
rldl_batch_update_settings rldl_batch_dims rldl_batch_last_times rldl_batch_stream_stats rldl_batch_counters rldl_batch_stream rldl_batch_select_kernel rldl_batch_set_tuning rldl_batch_lane_stats rldl_batch_get_trace rldl_batch_debug_sched rldl_batch_get_scaling rldl_batch_get_polish
rldl_batch_free
rldl_multi_setup rldl_multi_shards rldl_multi_shard_handle rldl_multi_solve rldl_multi_solve_to_host
rldl_multi_get_solution rldl_multi_get_info rldl_multi_update_bounds rldl_multi_partial_update_bounds
rldl_multi_update_lin_cost rldl_multi_warm_start rldl_multi_warm_start_shift rldl_multi_update_horizon
rldl_multi_update_rho rldl_multi_refactor rldl_multi_update_settings rldl_multi_free
rldl_device_count rldl_set_device rldl_version
""".split()

_lib = None


def lib():
    """Load librldl_b200.so (raises if it has not been built: there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build() / make -C osqp-recursive-ldl_b200")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.rldl_version.restype = C.c_char_p
    L.rldl_device_count.restype = c_int
    L.rldl_set_device.restype = c_int
    L.rldl_set_device.argtypes = [c_int]
    L.osqp_set_default_settings.argtypes = [C.POINTER(OSQPSettings)]
    L.osqp_set_default_settings.restype = None
    sig = {
        "osqp_setup": [C.POINTER(C.POINTER(OSQPWorkspace)), C.POINTER(OSQPData), C.POINTER(OSQPSettings)],
        "osqp_solve": [C.POINTER(OSQPWorkspace)],
        "osqp_solve_recursive": [C.POINTER(OSQPWorkspace)],
        "osqp_solve_combine_recursive": [C.POINTER(OSQPWorkspace)],
        "osqp_cleanup": [C.POINTER(OSQPWorkspace)],
        "osqp_update_lin_cost": [C.POINTER(OSQPWorkspace), vp],
        "osqp_update_bounds": [C.POINTER(OSQPWorkspace), vp, vp],
        "osqp_update_lower_bound": [C.POINTER(OSQPWorkspace), vp],
        "osqp_update_upper_bound": [C.POINTER(OSQPWorkspace), vp],
        "osqp_warm_start": [C.POINTER(OSQPWorkspace), vp, vp],
        "osqp_warm_start_x": [C.POINTER(OSQPWorkspace), vp],
        "osqp_warm_start_y": [C.POINTER(OSQPWorkspace), vp],
        "osqp_update_P": [C.POINTER(OSQPWorkspace), vp, vp, c_int],
        "osqp_update_A": [C.POINTER(OSQPWorkspace), vp, vp, c_int],
        "osqp_update_P_A": [C.POINTER(OSQPWorkspace), vp, vp, c_int, vp, vp, c_int],
        "osqp_update_rho": [C.POINTER(OSQPWorkspace), c_float],
        "osqp_setup_recursive": [C.POINTER(C.POINTER(OSQPWorkspace)), C.POINTER(OSQPDataRLDL),
                                 C.POINTER(OSQPSettings)] + [c_int] * 6,
        "osqp_setup_combine_recursive": [C.POINTER(C.POINTER(OSQPWorkspace)), C.POINTER(OSQPDataRLDL),
                                         C.POINTER(OSQPSettings)] + [c_int] * 7,
        "osqp_update_recursive": [C.POINTER(OSQPWorkspace), C.POINTER(OSQPDataRLDL), c_int],
        "osqp_update_Z_horizon": [C.POINTER(OSQPWorkspace), C.POINTER(OSQPDataRLDL), c_int],
        "osqp_update_X_horizon": [C.POINTER(OSQPWorkspace), C.POINTER(OSQPDataRLDL), c_int],
        "osqp_partial_update_bounds": [C.POINTER(OSQPWorkspace), c_int, c_int, vp, vp],
        "get_L_dimensions": [C.POINTER(OSQPWorkspace), C.POINTER(c_int), C.POINTER(c_int), C.POINTER(c_int)],
        "rldl_batch_setup": [C.POINTER(vp), C.POINTER(OSQPDataRLDL), C.POINTER(OSQPSettings)] + [c_int] * 7 + [vp] * 4,
        "rldl_batch_solve": [vp],
        "rldl_batch_solve_to_host": [vp] * 10,
        "rldl_batch_update_bounds": [vp, vp, vp],
        "rldl_batch_partial_update_bounds": [vp, c_int, c_int, vp, vp],
        "rldl_batch_update_lin_cost": [vp, vp],
        "rldl_batch_update_data_refactor": [vp, vp, vp, vp],
        "rldl_batch_warm_start": [vp, vp, vp],
        "rldl_batch_warm_start_shift": [vp],
        "rldl_batch_update_horizon": [vp, c_int],
        "rldl_batch_update_rho": [vp, vp],
        "rldl_batch_refactor": [vp],
        "rldl_batch_failed_factorisations": [vp, vp],
        "rldl_batch_update_model": [vp, vp],
        "rldl_batch_kkt_solve": [vp, vp],
        "rldl_batch_kkt_solve_static": [vp, vp, vp],
        "rldl_batch_get_solution": [vp, vp, vp],
        "rldl_batch_get_info": [vp] * 8,
        "rldl_batch_update_settings": [vp, C.POINTER(OSQPSettings)],
        "rldl_batch_get_scaling": [vp, vp, vp, vp],
        "rldl_batch_get_polish": [vp, vp],
        "rldl_batch_lane_stats": [vp, vp],
        "rldl_batch_get_trace": [vp, vp],
        "rldl_batch_dims": [vp] + [C.POINTER(c_int)] * 4,
        "rldl_batch_last_times": [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "rldl_batch_counters": [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(c_int)],
        "rldl_batch_stream_stats": [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "rldl_batch_free": [vp],
        "rldl_multi_setup": [C.POINTER(vp), vp, c_int, c_int, C.POINTER(OSQPDataRLDL), C.POINTER(OSQPSettings)] +
                            [c_int] * 7 + [vp] * 4,
        "rldl_multi_shards": [vp, vp, vp],
        "rldl_multi_solve": [vp],
        "rldl_multi_solve_to_host": [vp] * 10,
        "rldl_multi_get_solution": [vp, vp, vp],
        "rldl_multi_get_info": [vp] * 8,
        "rldl_multi_update_bounds": [vp, vp, vp],
        "rldl_multi_partial_update_bounds": [vp, c_int, c_int, vp, vp],
        "rldl_multi_update_lin_cost": [vp, vp],
        "rldl_multi_warm_start": [vp, vp, vp],
        "rldl_multi_warm_start_shift": [vp],
        "rldl_multi_update_horizon": [vp, c_int],
        "rldl_multi_update_rho": [vp, vp],
        "rldl_multi_refactor": [vp],
        "rldl_multi_update_settings": [vp, C.POINTER(OSQPSettings)],
        "rldl_multi_free": [vp],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = None if name in ("get_L_dimensions", "rldl_batch_dims", "rldl_batch_last_times",
                                      "rldl_batch_counters", "rldl_batch_stream_stats", "rldl_batch_free", "rldl_multi_free") else c_int
    L.rldl_batch_select_kernel.argtypes = [vp, c_int]
    L.rldl_batch_select_kernel.restype = c_int
    L.rldl_batch_set_tuning.argtypes = [vp, C.c_char_p, c_int]
    L.rldl_batch_set_tuning.restype = c_int
    L.rldl_multi_shard_handle.argtypes = [vp, c_int]
    L.rldl_multi_shard_handle.restype = vp
    L.rldl_batch_max_static_doubles.argtypes = []
    L.rldl_batch_max_static_doubles.restype = c_int
    L.rldl_batch_stream.argtypes = [vp]
    L.rldl_batch_stream.restype = vp
    for name in ("max_iter", "warm_start", "scaled_termination", "check_termination", "polish",
                 "polish_refine_iter", "verbose"):
        fn = getattr(L, "osqp_update_" + name); fn.argtypes = [C.POINTER(OSQPWorkspace), c_int]; fn.restype = c_int
    for name in ("eps_abs", "eps_rel", "eps_prim_inf", "eps_dual_inf", "alpha", "delta", "time_limit"):
        fn = getattr(L, "osqp_update_" + name); fn.argtypes = [C.POINTER(OSQPWorkspace), c_float]; fn.restype = c_int
    _lib = L
    return L


def default_settings(**kw) -> OSQPSettings:
    s = OSQPSettings()
    lib().osqp_set_default_settings(C.byref(s))
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_csc(t):
    """(m, n, p, i, x) numpy tuple -> csc struct borrowing the arrays (keep `t` alive)."""
    m, n, p, i, x = t
    M = csc()
    M.nzmax = max(1, len(x)); M.m = m; M.n = n; M.nz = -1
    M.p = p.ctypes.data_as(C.POINTER(c_int)); M.i = i.ctypes.data_as(C.POINTER(c_int))
    M.x = x.ctypes.data_as(C.POINTER(c_float))
    return M


class StageModel:
    """OSQPDataRLDL filled from an mpc.MPCModel (stage matrices kept alive here)."""

    def __init__(self, model):
        self.model = model
        self.tuples = model.stage_matrices()
        self.cscs = [make_csc(t) for t in self.tuples]
        d = OSQPDataRLDL()
        d.Q0, d.Qi, d.QN, d.A0, d.Ai, d.Aij, d.AN = [C.pointer(c) for c in self.cscs]
        d.nx, d.nu, d.ny, d.nt = model.nx, model.nu, model.ny, model.nt
        self.data = d


class Batch:
    """Many MPC QP instances sharing one stage model (rldl_batch_* in the C ABI)."""

    def __init__(self, model, N, Nmax, q, l, u, settings: OSQPSettings, rho=None):
        L = lib()
        self.sm = StageModel(model)
        self.batch = q.shape[0]
        self.nmax, self.mmax = q.shape[1], l.shape[1]
        q = np.ascontiguousarray(q, np.float64); l = np.ascontiguousarray(l, np.float64)
        u = np.ascontiguousarray(u, np.float64)
        rho = None if rho is None else np.ascontiguousarray(rho, np.float64)
        h = C.c_void_p()
        rc = L.rldl_batch_setup(C.byref(h), C.byref(self.sm.data), C.byref(settings), Nmax, N, model.nx,
                                model.nu, model.ny, model.nt, self.batch, _p(q), _p(l), _p(u), _p(rho))
        if rc:
            raise RuntimeError(f"rldl_batch_setup failed with code {rc}")
        self.h = h

    def dims(self):
        n, m, N, b = c_int(), c_int(), c_int(), c_int()
        lib().rldl_batch_dims(self.h, C.byref(n), C.byref(m), C.byref(N), C.byref(b))
        return n.value, m.value, N.value, b.value

    def solve(self):
        return lib().rldl_batch_solve(self.h)

    def solve_to_host(self, out):
        """rldl_batch_solve_to_host into the arrays of `out` (keys as get(); pinned arrays are written by the
        kernels while the solve runs, pageable ones by copies after it)."""
        return lib().rldl_batch_solve_to_host(self.h, *[_p(out[k]) for k in (
            "x", "y", "iter", "status", "obj", "pri_res", "dua_res", "rho_updates", "rho_estimate")])

    def get(self):
        B = self.batch
        out = dict(x=np.zeros((B, self.nmax)), y=np.zeros((B, self.mmax)), iter=np.zeros(B, np.int64),
                   status=np.zeros(B, np.int64), obj=np.zeros(B), pri_res=np.zeros(B),
                   dua_res=np.zeros(B), rho_updates=np.zeros(B, np.int64), rho_estimate=np.zeros(B))
        lib().rldl_batch_get_solution(self.h, _p(out["x"]), _p(out["y"]))
        lib().rldl_batch_get_info(self.h, *[_p(out[k]) for k in ("iter", "status", "obj", "pri_res",
                                                                  "dua_res", "rho_updates", "rho_estimate")])
        n, m, _, _ = self.dims()
        out["x"][:, n:] = 0.0
        out["y"][:, m:] = 0.0
        return out

    def update_bounds(self, l, u):
        return lib().rldl_batch_update_bounds(self.h, _p(np.ascontiguousarray(l)), _p(np.ascontiguousarray(u)))

    def partial_update_bounds(self, start, stop, l, u):
        return lib().rldl_batch_partial_update_bounds(self.h, start, stop, _p(np.ascontiguousarray(l)),
                                                      _p(np.ascontiguousarray(u)))

    def update_lin_cost(self, q):
        return lib().rldl_batch_update_lin_cost(self.h, _p(np.ascontiguousarray(q)))

    def update_data_refactor(self, q, l, u):
        return lib().rldl_batch_update_data_refactor(self.h, _p(np.ascontiguousarray(q)), _p(np.ascontiguousarray(l)),
                                                     _p(np.ascontiguousarray(u)))

    def warm_start(self, x=None, y=None):
        x = None if x is None else np.ascontiguousarray(x)
        y = None if y is None else np.ascontiguousarray(y)
        return lib().rldl_batch_warm_start(self.h, _p(x), _p(y))

    def warm_start_shift(self):
        return lib().rldl_batch_warm_start_shift(self.h)

    def update_horizon(self, N):
        return lib().rldl_batch_update_horizon(self.h, N)

    def update_rho(self, rho):
        return lib().rldl_batch_update_rho(self.h, _p(np.ascontiguousarray(rho, np.float64)))

    def refactor(self):
        return lib().rldl_batch_refactor(self.h)

    def update_model(self, mdl):
        """New values for the shared stage matrices (same shapes), then the full recursion."""
        sm = StageModel(mdl)
        self._models = getattr(self, "_models", []) + [sm]
        return lib().rldl_batch_update_model(self.h, C.byref(sm.data))

    def failed_factorisations(self):
        """(count, flag[batch]) — flag = 0, or 1 + the stage of the first non-positive pivot."""
        n, m, N, batch = self.dims()
        flag = np.zeros(batch, np.int64)
        return lib().rldl_batch_failed_factorisations(self.h, _p(flag)), flag

    def get_trace(self):
        t = np.zeros(self.batch)
        assert lib().rldl_batch_get_trace(self.h, _p(t)) == 0
        return t % 1e6, (t // 1e6).astype(int)

    def lane_stats(self):
        v = np.zeros(6, np.int64)
        assert lib().rldl_batch_lane_stats(self.h, _p(v)) == 0
        return v.tolist()

    def get_polish(self):
        sp = np.zeros(self.batch, np.int64)
        assert lib().rldl_batch_get_polish(self.h, _p(sp)) == 0
        return sp

    def get_scaling(self):
        """(D, E, c) of the Ruiz equilibration, None when scaling is off."""
        D = np.zeros((self.batch, self.nmax)); E = np.zeros((self.batch, self.mmax)); c = np.zeros(self.batch)
        if lib().rldl_batch_get_scaling(self.h, _p(D), _p(E), _p(c)):
            return None
        return D, E, c

    def update_settings(self, settings):
        return lib().rldl_batch_update_settings(self.h, C.byref(settings))

    def kkt_solve(self, rhs):
        rhs = np.ascontiguousarray(rhs, np.float64).copy()
        rc = lib().rldl_batch_kkt_solve(self.h, _p(rhs))
        if rc:
            raise RuntimeError(f"rldl_batch_kkt_solve failed with code {rc}")
        return rhs

    def kkt_solve_static(self, rhs):
        """rldl_batch_kkt_solve_static (no allocation on the call path); returns (rc, solution)."""
        rhs = np.ascontiguousarray(rhs, np.float64)
        sol = np.empty_like(rhs)
        return lib().rldl_batch_kkt_solve_static(self.h, _p(rhs), _p(sol)), sol

    def last_times(self):
        f, s = C.c_double(), C.c_double()
        lib().rldl_batch_last_times(self.h, C.byref(f), C.byref(s))
        return f.value, s.value

    def stream_stats(self):
        """(ms, instance-iterations) of the streaming ADMM kernel of the last solve alone."""
        t, it = C.c_double(), C.c_double()
        lib().rldl_batch_stream_stats(self.h, C.byref(t), C.byref(it))
        return t.value, it.value

    def counters(self):
        a, f, n = C.c_double(), C.c_double(), c_int()
        lib().rldl_batch_counters(self.h, C.byref(a), C.byref(f), C.byref(n))
        return a.value, f.value, n.value

    def select_kernel(self, which):
        return lib().rldl_batch_select_kernel(self.h, which)

    def set_tuning(self, **kw):
        """rldl_batch_set_tuning: lat_below, ring_depth, sched_s0, ctas_per_sm (-1 = default)."""
        for k, v in kw.items():
            rc = lib().rldl_batch_set_tuning(self.h, k.encode(), int(v))
            if rc:
                raise ValueError(f"rldl_batch_set_tuning({k!r}, {v}) refused")

    def stream(self):
        return lib().rldl_batch_stream(self.h)

    def free(self):
        if self.h:
            lib().rldl_batch_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Multi:
    """A batch sharded over devices inside the library (rldl_multi_* in the C ABI)."""

    def __init__(self, model, N, Nmax, q, l, u, settings: OSQPSettings, devices=None, nshards=0, rho=None):
        L = lib()
        self.sm = StageModel(model)
        self.batch, self.nmax, self.mmax = q.shape[0], q.shape[1], l.shape[1]
        q, l, u = (np.ascontiguousarray(v, np.float64) for v in (q, l, u))
        rho = None if rho is None else np.ascontiguousarray(rho, np.float64)
        dev = None if devices is None else np.ascontiguousarray(devices, np.int64)
        h = C.c_void_p()
        rc = L.rldl_multi_setup(C.byref(h), _p(dev), 0 if dev is None else len(dev), nshards, C.byref(self.sm.data),
                                C.byref(settings), Nmax, N, model.nx, model.nu, model.ny, model.nt, self.batch,
                                _p(q), _p(l), _p(u), _p(rho))
        if rc:
            raise RuntimeError(f"rldl_multi_setup failed with code {rc}")
        self.h = h

    def shards(self):
        n = lib().rldl_multi_shards(self.h, None, None)
        first, dev = np.zeros(n, np.int64), np.zeros(n, np.int64)
        lib().rldl_multi_shards(self.h, _p(first), _p(dev))
        return first.tolist(), dev.tolist()

    def empty_out(self):
        B = self.batch
        return dict(x=np.zeros((B, self.nmax)), y=np.zeros((B, self.mmax)), iter=np.zeros(B, np.int64),
                    status=np.zeros(B, np.int64), obj=np.zeros(B), pri_res=np.zeros(B), dua_res=np.zeros(B),
                    rho_updates=np.zeros(B, np.int64), rho_estimate=np.zeros(B))

    _KEYS = ("x", "y", "iter", "status", "obj", "pri_res", "dua_res", "rho_updates", "rho_estimate")

    def solve(self):
        return lib().rldl_multi_solve(self.h)

    def solve_to_host(self, out):
        return lib().rldl_multi_solve_to_host(self.h, *[_p(out[k]) for k in self._KEYS])

    def get(self):
        out = self.empty_out()
        assert lib().rldl_multi_get_solution(self.h, _p(out["x"]), _p(out["y"])) == 0
        assert lib().rldl_multi_get_info(self.h, *[_p(out[k]) for k in self._KEYS[2:]]) == 0
        return out

    def call(self, name, *arrays_or_ints):
        args = [_p(np.ascontiguousarray(a)) if isinstance(a, np.ndarray) else a for a in arrays_or_ints]
        return getattr(lib(), "rldl_multi_" + name)(self.h, *args)

    def free(self):
        if self.h:
            lib().rldl_multi_free(self.h)
            self.h = None
