"""rldl_multi_*: the batch sharded over devices INSIDE the library (one host thread and one handle per
shard, contiguous instance ranges, no collective; the gather is every shard writing its rows of the caller's arrays).
On a one-GPU box the shards share the device (concurrent handles); on a multi-GPU box they spread over all devices.
Checked against one RLDLBatch handle over the whole batch: identical status / iterations, x, y, obj within 1e-6 (which
ADMM kernel finishes an instance depends on the shard size: rounding-level differences)."""
import numpy as np
import pytest

import rldl_b200 as R
from rldl_b200 import mpc
from common import gpu_settings, problem, compare

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nshards", [0, 3, 8])
def test_multi_handle_matches_single_handle(nshards):
    ndev = int(R.lib().rldl_device_count())
    mdl, N, x0, q, l, u = problem("C1", 5000)
    one = R.Batch(mdl, N, N, q, l, u, gpu_settings(warm_start=1))
    assert one.solve() == 0
    want = one.get()
    n, m, _, _ = one.dims()
    mh = R.Multi(mdl, N, N, q, l, u, gpu_settings(warm_start=1), nshards=nshards)
    first, dev = mh.shards()
    S = nshards or ndev
    assert len(first) == S and first[0] == 0 and sorted(first) == first
    assert set(dev) == set(range(min(S, ndev)))                  # round-robin over every visible device
    out = mh.empty_out()
    assert mh.solve_to_host(out) == 0
    assert not compare(out, want, n, m)
    assert not compare(mh.get(), want, n, m)
    # closed-loop step through the multi handle: new x0 rows, shifted warm start, horizon shrink, new linear cost
    ny, nx = mdl.ny, mdl.nx
    dyn = np.ascontiguousarray(-(mpc.sample_x0(mdl, 5000, 77) @ mdl.A.T))
    q2 = q + 0.01
    for h, call in ((one, None), (mh, mh.call)):
        if call is None:
            assert h.partial_update_bounds(ny, ny + nx, dyn, dyn) == 0 and h.warm_start_shift() == 0
            assert h.update_horizon(N - 1) == 0 and h.update_lin_cost(q2) == 0 and h.solve() == 0
        else:
            assert call("partial_update_bounds", ny, ny + nx, dyn, dyn) == 0 and call("warm_start_shift") == 0
            assert call("update_horizon", N - 1) == 0 and call("update_lin_cost", q2) == 0 and h.solve() == 0
    n2, m2, _, _ = one.dims()
    assert not compare(mh.get(), one.get(), n2, m2)
    # an invalid update is refused by the shard that holds the offending instance
    lbad = l.copy()
    lbad[4321, 5] = u[4321, 5] + 1.0                    # an input row (finite bounds)
    assert mh.call("update_bounds", lbad, u) == 1
    one.free(); mh.free()
