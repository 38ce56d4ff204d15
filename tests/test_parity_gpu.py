"""GPU parity tests: the CUDA path (through the C ABI) against the unmodified
reference (oracle/_ref, run in a child process) on identical seeded inputs."""
import numpy as np
import pytest

import rldl_b200 as R
from common import gpu_settings, ref_run, problem, compare, rel_err

pytestmark = pytest.mark.gpu


def mpc_x0(mdl, batch, seed):
    from rldl_b200 import mpc
    return mpc.sample_x0(mdl, batch, seed)


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


def test_scheduler_does_not_change_results(monkeypatch):
    """The device priority scheduler only reorders / migrates instances: with a batch large enough to
    oversubscribe the groups and force preemptions, results are bitwise those of first-come order."""
    mdl, N, x0, q, l, u = problem("C1", 40000)
    outs = []
    for s0 in (0, 25):
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
        # streaming kernel only: which instances the latency-mode kernel finishes is timing dependent
        # (rounding-level differences)
        gb.set_tuning(lat_below=0, sched_s0=s0)
        assert gb.solve() == 0
        outs.append(gb.get())
        gb.free()
    assert (outs[0]["iter"] == outs[1]["iter"]).all() and (outs[0]["status"] == outs[1]["status"]).all()
    assert np.array_equal(outs[0]["x"], outs[1]["x"]) and np.array_equal(outs[0]["y"], outs[1]["y"])
    assert outs[0]["iter"].max() > outs[0]["iter"].min()  # the batch really is imbalanced


@pytest.mark.parametrize("below", [0, 16, 100000])
@pytest.mark.parametrize("name,batch", [("C4", 48), ("C1", 64)])
def test_adaptive_rho_streaming_matches_generic_kernel(name, batch, below, monkeypatch):
    """Adaptive rho: the streaming / latency-mode kernels park an instance, the host re-factorises the parked
    ones in one batched launch and resumes them; the generic kernel re-factorises in place.  Same iteration,
    whether the streaming kernel runs alone (0), hands its tail to the latency-mode kernel (16) or the
    latency-mode kernel runs alone (100000)."""
    mdl, N, x0, q, l, u = problem(name, batch)
    rng = np.random.default_rng(5)
    rho = rng.choice([1e-3, 0.1, 10.0], size=batch)
    outs = []
    for which in (0, 1):
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings(adaptive_rho=1, adaptive_rho_interval=25,
                                                      adaptive_rho_tolerance=1.5), rho=rho)
        assert gb.select_kernel(which) == 0
        gb.set_tuning(lat_below=below)
        assert gb.solve() == 0
        outs.append(gb.get())
        n, m, _, _ = gb.dims()
        gb.free()
    assert outs[0]["rho_updates"].sum() > 0
    assert (outs[0]["rho_updates"] == outs[1]["rho_updates"]).all()
    bad = compare(outs[0], outs[1], n, m)
    assert not bad, bad[:5]
    # rho_estimate is a ratio of converged residuals: only its magnitude is comparable, and only where the
    # residuals are above the cancellation noise of their own evaluation (a residual of 1e-10 on O(1) data
    # differs by a factor between two summation orders)
    live = np.minimum(np.minimum(outs[0]["pri_res"], outs[0]["dua_res"]),
                      np.minimum(outs[1]["pri_res"], outs[1]["dua_res"])) > 1e-8
    assert np.allclose(outs[0]["rho_estimate"][live], outs[1]["rho_estimate"][live], rtol=0.5)


def test_batch_update_bounds_rejects_l_above_u_and_keeps_problem():
    """osqp_update_bounds semantics (osqp.c:793-801) for the batch entry point: validated on the device,
    invalid input returns 1 and leaves the resident problem untouched."""
    mdl, N, x0, q, l, u = problem("C1", 16)
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    assert gb.solve() == 0
    ref = gb.get()
    lbad = l.copy()
    lbad[3, 7] = u[3, 7] + 1.0
    assert gb.update_bounds(lbad, u) == 1
    assert gb.solve() == 0
    out = gb.get()
    assert (out["iter"] == ref["iter"]).all() and np.array_equal(out["x"], ref["x"])
    assert gb.update_bounds(l, u) == 0
    gb.free()


@pytest.mark.parametrize("name,batch", [("C1", 300), ("C3", 200), ("C4", 160)])
@pytest.mark.parametrize("mode", ["latency_only", "handoff"])
def test_latency_mode_kernel_matches_streaming_kernel(name, batch, mode, monkeypatch):
    """The CTA-per-instance latency-mode kernel (re-associated block substitution, factor resident in shared
    memory) alone, and as the finisher of a streaming launch that suspends its last instances, against the
    streaming kernel alone: same status / iterations, x, y, obj within 1e-6."""
    mdl, N, x0, q, l, u = problem(name, batch)
    outs = []
    for below in (0, 100000 if mode == "latency_only" else batch // 3):
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
        gb.set_tuning(lat_below=below)
        assert gb.solve() == 0
        outs.append(gb.get())
        n, m, _, _ = gb.dims()
        gb.free()
    bad = compare(outs[1], outs[0], n, m)
    assert not bad, bad[:5]


@pytest.mark.parametrize("nx,nu", [(2, 1), (4, 1), (6, 2), (6, 3), (8, 4), (16, 4)])
def test_other_specialised_shapes_match_reference(nx, nu, monkeypatch):
    """The shape table of rldl_stream_common.cuh beyond the BASELINE shapes (odd and even block sizes in each
    lane-group width): the recursion kernel (one KKT solve), the streaming kernel alone and the latency-mode
    kernel alone, each against the unmodified reference."""
    from rldl_b200 import mpc
    N, batch = 9, 24
    mdl = mpc.random_model(nx, nu, seed=100 + nx * 10 + nu)
    x0 = mpc.sample_x0(mdl, batch, 77)
    q, l, u = mdl.vectors(x0, N, N + 1)
    rng = np.random.default_rng(3)
    res = ref_run(mdl, N, N + 1, q, l, u, [("solve",), ("get",), ("dims",)])
    ref, (n, m) = res[1], res[2]
    rhs = rng.standard_normal((batch, n + m))
    kk = ref_run(mdl, N, N + 1, q, l, u, [("kkt_solve", k, rhs[k]) for k in range(3)])
    for below in (0, 100000):
        gb = R.Batch(mdl, N, N + 1, q, l, u, gpu_settings())
        gb.set_tuning(lat_below=below)
        assert gb.select_kernel(2) == 0          # the shape-specialised kernels serve this shape
        got = gb.kkt_solve(rhs)
        for k in range(3):
            assert rel_err(got[k], kk[k]) < 1e-9, (nx, nu, k)
        assert gb.solve() == 0
        bad = compare(gb.get(), ref, n, m)
        assert not bad, (nx, nu, below, bad[:5])
        gb.free()


@pytest.mark.parametrize("nx,nu,N", [(12, 4, 2), (12, 4, 7), (12, 4, 12), (12, 4, 22), (6, 2, 13), (16, 4, 10), (2, 1, 6)])
def test_latency_mode_block_tree_levels_match_reference(nx, nu, N):
    """The latency-mode kernel with 1..5 levels of the block tree under its sequential chain (tuning knob
    "lat_levels": 1 = pairs of stages, the default; deeper trees shorten the chain and add hops behind it), over
    horizons with every shape of leftover (N - 1 = 1, 6, 11, 21, ...), with lockstep items ("lat_lockstep") and
    without: each against the unmodified reference."""
    from rldl_b200 import mpc
    batch = 12
    mdl = mpc.random_model(nx, nu, seed=500 + nx * 10 + nu)
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, 11 + N), N, N)
    res = ref_run(mdl, N, N, q, l, u, [("solve",), ("get",), ("dims",)])
    ref, (n, m) = res[1], res[2]
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    for lockstep in (1, 0):     # a warp per item (two lane groups per warp only) / a lane group per item
        for lv in (1, 2, 3, 4, 5):
            gb.set_tuning(lat_below=100000, lat_levels=lv, lat_lockstep=lockstep)
            assert gb.solve() == 0
            bad = compare(gb.get(), ref, n, m)
            assert not bad, (nx, nu, N, lockstep, lv, bad[:5])
    gb.free()


@pytest.mark.parametrize("name,batch,below,kernel", [("C1", 3000, 1024, 0), ("C1", 500, 100000, 0), ("C3", 300, 100, 0),
                                                     ("C1", 64, 0, 1)],
                         ids=["stream+tail", "latency", "C3 hand-off", "generic"])
@pytest.mark.parametrize("pinned", [True, False])
def test_solve_to_host_delivers_what_get_returns(name, batch, below, kernel, pinned):
    """rldl_batch_solve_to_host: the kernels write every finished instance to PINNED host arrays themselves
    (pageable arrays: copies after the solve); bitwise what rldl_batch_get_solution / _get_info return, including
    OSQP_NAN rows for an instance without a solution."""
    import torch
    mdl, N, x0, q, l, u = problem(name, batch)
    b = mdl.nx + mdl.ny
    l[1, b:b + mdl.nx] = 10.0; u[1, b:b + mdl.nx] = 11.0      # instance 1: primal infeasible
    gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    assert gb.select_kernel(kernel) == 0
    gb.set_tuning(lat_below=below)
    shapes = dict(x=q.shape, y=l.shape, iter=batch, status=batch, obj=batch, pri_res=batch, dua_res=batch,
                  rho_updates=batch, rho_estimate=batch)
    ints = ("iter", "status", "rho_updates")
    tens = {k: torch.full(shp if isinstance(shp, tuple) else (shp,), -77, dtype=torch.int64 if k in ints else torch.float64)
            for k, shp in shapes.items()}
    if pinned:
        tens = {k: t.pin_memory() for k, t in tens.items()}
    out = {k: t.numpy() for k, t in tens.items()}
    assert gb.solve_to_host(out) == 0
    want = gb.get()
    n, m, _, _ = gb.dims()
    gb.free()
    assert want["status"][1] in (-3, 3, -2) and (want["status"] == 1).sum() >= batch - 2
    assert name != "C1" or want["status"][1] == -3
    for k in ("iter", "status", "obj", "pri_res", "dua_res", "rho_updates", "rho_estimate"):
        assert np.array_equal(out[k], want[k]), k
    assert np.array_equal(out["x"][:, :n], want["x"][:, :n]) and np.array_equal(out["y"][:, :m], want["y"][:, :m])
    if want["status"][1] in (-3, 3):
        assert (out["x"][1, :n] == float(0x7fc00000)).all()


def test_lane_first_use_in_a_fresh_process_does_not_deadlock():
    """The concurrent latency-mode CTAs spin until a kernel launched after them (k_set_tail_flag) has run; with CUDA's
    lazy module loading the FIRST launch of that kernel in a process used to wait for the spinning CTAs (adaptive rho:
    parked instances keep the unfinished count above zero, so they never left) — a deadlock that only shows when the
    lane is the first thing a process does.  tools/exp_hang.py runs exactly that, in a fresh interpreter."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cp = subprocess.run([sys.executable, os.path.join(root, "tools", "exp_hang.py"), "8", "1"], capture_output=True,
                        text=True, timeout=150)
    assert cp.returncode == 0 and "ALL DONE" in cp.stdout, (cp.returncode, cp.stdout[-800:], cp.stderr[-800:])


@pytest.mark.parametrize("name,batch", [("C1", 3000), ("C3", 500), ("C1", 20)])
def test_update_data_refactor_equals_the_three_calls(name, batch):
    """rldl_batch_update_data_refactor (chunked upload pipelined against typing / records / recursion) ==
    rldl_batch_update_bounds + rldl_batch_update_lin_cost + rldl_batch_refactor; a refused update keeps the bounds."""
    import torch
    mdl, N, x0, q, l, u = problem(name, batch)
    _, l2, u2 = mdl.vectors(mpc_x0(mdl, batch, 55), N)
    q2 = q + 0.05 * np.random.default_rng(4).standard_normal(q.shape)
    b = mdl.nx + mdl.ny
    l2[:, b:b + mdl.nx] = -1e30; u2[:, b:b + mdl.nx] = 1e30      # stage-1 state box loose: row classes change
    a = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    c = R.Batch(mdl, N, N, q, l, u, gpu_settings())
    assert a.solve() == 0 and c.solve() == 0
    assert a.update_bounds(l2, u2) == 0 and a.update_lin_cost(q2) == 0 and a.refactor() == 0 and a.solve() == 0
    pin = lambda v: torch.from_numpy(np.ascontiguousarray(v)).pin_memory().numpy()
    assert c.update_data_refactor(pin(q2), pin(l2), pin(u2)) == 0
    assert (c.get()["status"] == -10).all()                        # info reset, like the reference's updates
    assert c.solve() == 0
    n, m, _, _ = a.dims()
    wa, wc = a.get(), c.get()
    assert (wa["iter"] == wc["iter"]).all() and (wa["status"] == wc["status"]).all()
    assert np.abs(wa["x"] - wc["x"]).max() < 1e-9 and np.abs(wa["y"] - wc["y"]).max() < 1e-9
    lbad = l2.copy()
    lbad[batch // 2, mdl.ny - 1] = u2[batch // 2, mdl.ny - 1] + 1.0    # an input row of stage 0 (finite bounds)
    assert c.update_data_refactor(q2, lbad, u2) == 1
    assert c.solve() == 0
    wc2 = c.get()
    assert (wc2["iter"] == wc["iter"]).all() and np.abs(wc2["x"] - wc["x"]).max() < 1e-9
    a.free(); c.free()
