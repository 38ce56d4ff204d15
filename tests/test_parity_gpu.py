"""GPU parity tests: the CUDA path (through the C ABI) against the unmodified
reference (oracle/_ref, run in a child process) on identical seeded inputs."""
import numpy as np
import pytest

import rldl_b200 as R
from common import gpu_settings, ref_run, problem, compare, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,batch", [("C1", 8), ("C3", 6), ("C4", 6)])
def test_kkt_solve_matches_reference(name, batch):
    """One KKT solve incl. the z_tilde step (qdldl_interface.c:559-585) on random rhs."""
    mdl, N, x0, q, l, u = problem(name, batch)
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    n, m, _, _ = gb.dims()
    rng = np.random.default_rng(7)
    rhs = rng.standard_normal((batch, n + m))
    got = gb.kkt_solve(rhs)
    want = ref_run(mdl, N, N, q, l, u, [("kkt_solve", k, rhs[k]) for k in range(batch)])
    for k in range(batch):
        assert rel_err(got[k], want[k]) < 1e-9, (name, k, rel_err(got[k], want[k]))
    gb.free()


@pytest.mark.parametrize("name,batch", [("C1", 32), ("C3", 16), ("C4", 16)])
def test_cold_solve_matches_reference(name, batch):
    mdl, N, x0, q, l, u = problem(name, batch)
    ref = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",)])[1]
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    assert gb.solve() == 0
    out = gb.get()
    n, m, _, _ = gb.dims()
    bad = compare(out, ref, n, m)
    assert not bad, bad[:5]
    gb.free()


@pytest.mark.parametrize("name,batch", [("C1", 70), ("C3", 40), ("C4", 24)])
def test_streaming_kernel_matches_generic_kernel(name, batch):
    """The TMA-streaming ADMM kernel and the generic kernel run the same iteration."""
    mdl, N, x0, q, l, u = problem(name, batch)
    outs = []
    for which in (0, 1):
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
        assert gb.select_kernel(which) == 0
        assert gb.solve() == 0
        outs.append(gb.get())
        n, m, _, _ = gb.dims()
        gb.free()
    bad = compare(outs[0], outs[1], n, m)
    assert not bad, bad[:5]
