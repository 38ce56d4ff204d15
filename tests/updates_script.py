"""The op sequence of tests/golden/updates_C1.npz (make_golden.updates) replayed on any batch facade
(oracle.OracleBatch or rldl_b200.Batch) — shared by the CPU oracle test and the GPU parity test."""
import numpy as np

from common import rel_err, REL_TOL
from rldl_b200 import mpc


def inputs(g, Nmax_pad=1):
    cfg = mpc.CONFIGS["C1"]
    mdl, N = mpc.make_model(cfg), cfg["N"]
    x0 = mpc.sample_x0(mdl, int(g["batch"]), int(g["seed"]))
    q, l, u = mdl.vectors(x0, N, N + Nmax_pad)
    return mdl, N, q, l, u


def check(out, g, tag, n, m, take):
    assert (out["status"] == g[f"status_{tag}"][:take]).all(), (tag, out["status"])
    assert (out["iter"] == g[f"iter_{tag}"][:take]).all(), (tag, out["iter"], g[f"iter_{tag}"][:take])
    for k in range(take):
        assert rel_err(out["x"][k, :n], g[f"x_{tag}"][k]) < REL_TOL, (tag, k)
        assert rel_err(out["y"][k, :m], g[f"y_{tag}"][k]) < REL_TOL, (tag, k)
        assert abs(out["obj"][k] - g[f"obj_{tag}"][k]) <= REL_TOL * abs(g[f"obj_{tag}"][k]), (tag, k)


def replay(b, g, take):
    """b was built on inputs(g) (first `take` instances) with warm_start=1."""
    n, m = int(g["n"]), int(g["m"])
    s, e = int(g["s"]), int(g["e"])
    t = slice(0, take)
    assert b.solve() == 0
    check(b.get(), g, 0, n, m, take)
    assert b.update_lin_cost(g["q2"][t]) == 0
    assert b.partial_update_bounds(s, e, np.ascontiguousarray(g["l2"][t, s:e]),
                                   np.ascontiguousarray(g["u2"][t, s:e])) == 0
    assert b.solve() == 0
    check(b.get(), g, 1, n, m, take)
    assert b.warm_start(g["xw"][t], g["yw"][t]) == 0
    assert b.solve() == 0
    check(b.get(), g, 2, n, m, take)
    assert b.update_bounds(g["lt"][t], g["ut"][t]) == 0
    assert b.warm_start(g["xw"][t], None) == 0
    assert b.solve() == 0
    check(b.get(), g, 3, n, m, take)
