"""Polish (settings.polish = 1, reference src/polish.c:212-350) on the B200 path — SURVEY §8f-4.

Oracle: the reference's stock path with polish on (golden tests/golden/polish_*.npz; accepted, rejected and
not-attempted cases), and live runs for the streaming shapes.  The device runs the reduced KKT of the guessed
active set through the same stage recursion (sigma = delta, rho = 1/delta on active rows, inactive rows removed),
iterative refinement, the normal-cone projection and the acceptance test (k_polish).  Bar: status_polish, status,
iteration count equal; x, y, obj within 1e-6 relative (the polished residuals themselves are rounding noise).
"""
import os
import sys

import numpy as np
import pytest

import rldl_b200 as R
from common import gpu_settings, ref_run, compare, problem

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
sys.path.insert(0, GOLD)


@pytest.mark.parametrize("name", ["C1", "C3", "C4"])
@pytest.mark.parametrize("kernel", [0, 1], ids=["auto", "generic"])
def test_polish_matches_golden(name, kernel):
    from make_golden import POLISH
    g = np.load(os.path.join(GOLD, f"polish_{name}.npz"))
    batch, prob, kw = POLISH[name]
    mdl, N, x0, q, l, u = prob(name, batch)
    n, m = int(g["n"]), int(g["m"])
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(**kw))
    assert gb.select_kernel(kernel) == 0
    assert gb.solve() == 0
    out = gb.get()
    sp = gb.get_polish()
    gb.free()
    assert (sp == g["status_polish"]).all(), (sp, g["status_polish"])
    ref = dict(status=g["status"], iter=g["iter"], obj=g["obj"],
               x=np.pad(g["x"], ((0, 0), (0, q.shape[1] - n))), y=np.pad(g["y"], ((0, 0), (0, l.shape[1] - m))))
    bad = compare(out, ref, n, m)
    assert not bad, bad[:3]
    acc = sp == 1
    assert (out["pri_res"][acc] < 1e-8).all() and (out["dua_res"][acc] < 1e-8).all()
    # rejected / not attempted: the ADMM residuals stay
    assert np.allclose(out["pri_res"][~acc], g["pri_res"][~acc], rtol=1e-6, atol=1e-12)


def test_polish_after_streaming_solve_matches_reference():
    """The default kernels (streaming / latency mode) solve, k_polish refines: C5-shaped batch with the bench
    tolerances against the reference's stock path; a second, warm-started solve starts from the polished point."""
    mdl, N, x0, q, l, u = problem("C3", 40)
    kw = dict(polish=1, warm_start=1)
    ref = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",), ("get_polish",), ("solve",), ("get",)], path="stock", **kw)
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(**kw))
    assert gb.select_kernel(2) == 0
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    assert (gb.get_polish() == ref[2]).all()
    bad = compare(out, ref[1], n, m)
    assert not bad, bad[:3]
    assert gb.solve() == 0
    bad = compare(gb.get(), ref[4], n, m)
    gb.free()
    assert not bad, bad[:3]


def test_scalar_api_polish():
    """osqp_update_polish(1) + osqp_solve_recursive: info->status_polish, polished obj_val (reference osqp.h:363)."""
    from test_api_gpu import Scalar
    mdl, N, x0, q, l, u = problem("C1", 1, Nmax=21)
    sc = Scalar(mdl, N, 21, q[0], l[0], u[0], gpu_settings(eps_abs=1e-3, eps_rel=1e-3))
    assert sc.rc == 0 and sc.solve() == 0
    assert sc.info.status_polish == 0
    rough = sc.out()
    assert sc.L.osqp_update_polish(sc.w, 1) == 0
    assert sc.solve() == 0
    assert sc.info.status_val == 1 and sc.info.status_polish == 1
    assert sc.info.pri_res < 1e-9 and sc.info.dua_res < 1e-9 and sc.info.polish_time > 0
    qs, ls, us = mdl.vectors(x0, N)
    ref = ref_run(mdl, N, N, qs, ls, us, [("solve",), ("get",)], path="stock", eps_abs=1e-3, eps_rel=1e-3, polish=1)[1]
    n, m = sc.dims()
    bad = compare(sc.out(), ref, n, m)
    assert not bad, bad
    assert np.abs(rough["x"] - sc.out()["x"]).max() > 1e-6      # the polished point differs from the ADMM iterate
    sc.free()
