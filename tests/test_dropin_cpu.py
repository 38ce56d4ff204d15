"""Drop-in boundary, CPU side: the user program of tests/dropin/mpc_user.c (written against the
reference's own headers: include/osqp.h, include/recursive_ldl.h, include/cs.h).

* linked against the unmodified reference it reproduces the committed golden output
  (tests/golden/dropin_mpc_user_ref.txt, made by that same binary), which pins the fixture the GPU
  test compares the product against;
* linked against librldl_b200.so — the SAME object file — it starts, and without a CUDA device
  osqp_setup_recursive fails loudly with OSQP_LINSYS_SOLVER_LOAD_ERROR (there is no CPU path);
* the CSC helpers of the reference's cs.h exported by the product library behave as documented.
"""
import ctypes as C
import os

import numpy as np
import pytest

import rldl_b200 as R
from dropin_common import GOLDEN, GOLDEN_API, assert_same_run, binary, parse, run


def _need(name):
    if not os.path.exists(binary(name)):
        pytest.skip(f"{name} not built (make -C tests/dropin needs /root/reference)")


def test_reference_linked_user_program_reproduces_the_golden_output():
    _need("dropin_ref")
    cp = run("dropin_ref")
    assert cp.returncode == 0, cp.stderr[-2000:]
    got, want = parse(cp.stdout), parse(open(GOLDEN).read())
    assert want["rc"] == {"SETUP": 0, "BADBOUNDS": 1, "DONE": 0} and want["dims"] == (324, 324)
    assert [want["solves"][t]["iter"] for t in want["solves"]] == [175, 125, 75, 75, 50, 50, 175]
    assert_same_run(got, want, 1e-9, "reference vs golden")


def test_reference_linked_api_program_reproduces_the_golden_output():
    _need("api_ref")
    cp = run("api_ref")
    assert cp.returncode == 0, cp.stderr[-2000:]
    got, want = parse(cp.stdout), parse(open(GOLDEN_API).read())
    assert want["rc"] == {"SETUP generic": 0, "BADRHO": 1, "CLEANUP": 0, "SETUP combine": 0, "SETUP unconstrained": 0,
                          "SETUP non_cvx_small_sigma": 5, "SETUP non_cvx": 0, "SETUP infeasibility1": 0,
                          "SETUP infeasibility2": 0, "SETUP infeasibility3": 0, "SETUP infeasibility4": 0,
                          "SETUP defaults": 0, "SETUP badly_scaled": 0, "DONE": 0}
    # part D: default settings (Ruiz scaling 10, adaptive rho) + polish; the reference keeps a stale status_polish
    assert want["solves"]["defaults_polish"]["status_polish"] == 1
    assert abs(want["solves"]["defaults_polish"]["obj"] - 1.88) < 1e-12            # polished to the known answer
    assert want["solves"]["badly_scaled_bounds_polish"]["status_polish"] == 1
    sv = {t: r["status_val"] for t, r in want["solves"].items()}
    assert (sv["unconstrained"], sv["non_cvx"], sv["infeasibility1"], sv["infeasibility2"], sv["infeasibility3"]) == \
        (1, -7, 1, -3, -4)
    assert abs(want["solves"]["unconstrained"]["obj"] + 19.209752026813277) < 1e-5   # tests/unconstrained
    assert abs(want["solves"]["infeasibility1"]["obj"] + 1.5) < 1e-3                 # tests/primal_dual_infeasibility
    assert abs(want["solves"]["qp"]["obj"] - 1.88) < 1e-5          # the reference's basic_qp known answer
    assert_same_run(got, want, 1e-9, "reference vs golden (api_user)")


@pytest.mark.parametrize("name", ["dropin_b200", "dropin_b200_ownhdr", "api_b200"])
def test_product_linked_user_program_fails_loudly_without_a_device(name):
    _need(name)
    if R.lib().rldl_device_count() > 0:
        pytest.skip("a CUDA device is present; covered by the gpu test")
    cp = run(name)
    assert cp.returncode == 13, (cp.returncode, cp.stderr[-500:])      # 10 + OSQP_LINSYS_SOLVER_LOAD_ERROR
    assert "rc 3" in cp.stdout.splitlines()[0] and cp.stdout.startswith("SETUP")
    assert "no CPU path" in cp.stderr


class _csc(C.Structure):
    _fields_ = [("nzmax", C.c_longlong), ("m", C.c_longlong), ("n", C.c_longlong), ("p", C.POINTER(C.c_longlong)),
                ("i", C.POINTER(C.c_longlong)), ("x", C.POINTER(C.c_double)), ("nz", C.c_longlong)]


def test_csc_helpers_of_the_reference_cs_h():
    lib = R.lib()
    P = C.POINTER(_csc)
    lib.csc_matrix.restype = P
    lib.csc_matrix.argtypes = [C.c_longlong] * 3 + [C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    lib.csc_spalloc.restype = P
    lib.csc_spalloc.argtypes = [C.c_longlong] * 5
    lib.csc_spfree.argtypes = [P]
    lib.copy_csc_mat.restype = P
    lib.copy_csc_mat.argtypes = [P]
    lib.csc_to_triu.restype = P
    lib.csc_to_triu.argtypes = [P]
    lib.triplet_to_csc.restype = P
    lib.triplet_to_csc.argtypes = [P, C.POINTER(C.c_longlong)]
    lib.osqp_version.restype = C.c_char_p
    assert lib.osqp_version() == b"0.6.0"

    # a full symmetric 3x3 in triplet form, entries in scrambled order (cs.h:92)
    dense = np.array([[4.0, 1.0, 0.0], [1.0, 5.0, 2.0], [0.0, 2.0, 6.0]])
    ent = [(r, c) for c in range(3) for r in range(3) if dense[r, c] != 0.0]
    ent = ent[::2] + ent[1::2]
    T = lib.csc_spalloc(3, 3, len(ent), 1, 1)
    assert T.contents.nz == 0 and T.contents.nzmax == len(ent)
    for k, (r, c) in enumerate(ent):
        T.contents.i[k], T.contents.p[k], T.contents.x[k] = r, c, dense[r, c]
    T.contents.nz = len(ent)
    where = (C.c_longlong * len(ent))()
    M = lib.triplet_to_csc(T, where)
    assert M.contents.nz == -1 and M.contents.p[3] == len(ent)
    back = np.zeros((3, 3))
    for c in range(3):
        for e in range(M.contents.p[c], M.contents.p[c + 1]):
            back[M.contents.i[e], c] += M.contents.x[e]
    assert (back == dense).all()
    for k, (r, c) in enumerate(ent):
        assert M.contents.i[where[k]] == r and M.contents.x[where[k]] == dense[r, c]

    U = lib.csc_to_triu(M)                                              # cs.h:113
    triu = np.zeros((3, 3))
    for c in range(3):
        for e in range(U.contents.p[c], U.contents.p[c + 1]):
            assert U.contents.i[e] <= c
            triu[U.contents.i[e], c] = U.contents.x[e]
    assert (triu == np.triu(dense)).all() and U.contents.p[3] == 5

    K = lib.copy_csc_mat(U)                                             # cs.h:70
    assert K.contents.p[3] == 5 and [K.contents.x[e] for e in range(5)] == [U.contents.x[e] for e in range(5)]
    assert C.addressof(K.contents.x.contents) != C.addressof(U.contents.x.contents)

    rect = lib.csc_spalloc(2, 3, 1, 1, 0)                               # not square -> NULL (cs.c)
    rect.contents.p[0] = rect.contents.p[1] = rect.contents.p[2] = rect.contents.p[3] = 0
    assert not lib.csc_to_triu(rect)

    x = (C.c_double * 2)(1.0, 2.0); i = (C.c_longlong * 2)(0, 1); p = (C.c_longlong * 3)(0, 1, 2)
    W = lib.csc_matrix(2, 2, 2, x, i, p)                                # cs.h:26: wraps, no copy
    assert W.contents.nz == -1 and C.addressof(W.contents.x.contents) == C.addressof(x)
    for h in (T, M, U, K, rect):
        lib.csc_spfree(h)
