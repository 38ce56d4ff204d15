"""GPU parity tests: the CUDA path (through the C ABI) against the unmodified
reference (oracle/_ref) on identical seeded inputs."""
import numpy as np
import pytest

import rldl_b200 as R
from oracle import ref_py
from common import settings_pair, problem, compare, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,batch", [("C1", 8), ("C3", 6), ("C4", 6)])
def test_kkt_solve_matches_reference(name, batch):
    """One KKT solve incl. the z_tilde step (qdldl_interface.c:559-585) on random rhs."""
    mdl, N, x0, q, l, u = problem(name, batch)
    sg, sr = settings_pair()
    rb = ref_py.RefBatch(mdl, N, N, q, l, u, sr, path="recursive")
    gb = R.Batch(mdl, N, N, q, l, u, sg)
    n, m, _, _ = gb.dims()
    rng = np.random.default_rng(7)
    rhs = rng.standard_normal((batch, n + m))
    got = gb.kkt_solve(rhs)
    for k in range(batch):
        want = rb.kkt_solve(k, rhs[k])
        assert rel_err(got[k], want) < 1e-9, (name, k, rel_err(got[k], want))
    gb.free(); rb.free()


@pytest.mark.parametrize("name,batch", [("C1", 32), ("C3", 16), ("C4", 16)])
def test_cold_solve_matches_reference(name, batch):
    mdl, N, x0, q, l, u = problem(name, batch)
    sg, sr = settings_pair()
    rb = ref_py.RefBatch(mdl, N, N, q, l, u, sr, path="recursive")
    rb.solve(); ref = rb.get()
    gb = R.Batch(mdl, N, N, q, l, u, sg)
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    bad = compare(out, ref, n, m)
    assert not bad, bad[:5]
    gb.free(); rb.free()
