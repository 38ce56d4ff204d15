"""rldl_batch_kkt_solve_static — the allocation-free KKT solve added (no
counterpart in the reference): right-hand sides staged in a fixed file-scope device array instead of a
cudaMalloc'ed buffer.  Checked through the C ABI against the existing entry point rldl_batch_kkt_solve on two calls
in a row (the array keeps the previous call's contents), at the tolerance the KKT-solve tests use (1e-9 relative);
sizes beyond the fixed scratch are refused.  Collected last on purpose."""
import numpy as np
import pytest

import rldl_b200 as R
from common import gpu_settings, problem, rel_err

pytestmark = pytest.mark.gpu


def test_static_kkt_solve_matches_the_allocating_entry_point_twice():
    limit = int(R.lib().rldl_batch_max_static_doubles())
    assert limit == 512 * 1024
    for name, batch in (("C1", 48), ("C3", 20)):
        mdl, N, x0, q, l, u = problem(name, batch)
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
        n, m, _, _ = gb.dims()
        assert batch * (n + m) <= limit
        rng = np.random.default_rng(17)
        for call in range(2):                                  # second call: the scratch holds the first call's data
            rhs = rng.standard_normal((batch, n + m)) * (1.0 + 9.0 * call)
            want = gb.kkt_solve(rhs)                           # existing entry point (cudaMalloc'ed staging buffer)
            rc, got = gb.kkt_solve_static(rhs)
            assert rc == 0
            assert np.isfinite(got).all()
            assert rel_err(got, want) < 1e-9, (name, call, rel_err(got, want))
            for k in range(0, batch, 7):
                assert rel_err(got[k], want[k]) < 1e-9, (name, call, k)
        gb.free()


def test_static_kkt_solve_refuses_sizes_beyond_its_scratch():
    limit = int(R.lib().rldl_batch_max_static_doubles())
    mdl, N, x0, q, l, u = problem("C1", 1700)                  # 1700 * 324 = 550 800 doubles > 524 288
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    n, m, _, _ = gb.dims()
    assert 1700 * (n + m) > limit
    rhs = np.ones((1700, n + m))
    rc, got = gb.kkt_solve_static(rhs)
    assert rc == 1
    want = gb.kkt_solve(rhs)                                   # the allocating entry point still serves the size
    assert np.isfinite(want).all()
    gb.free()
