"""Scenario drivers shared by every implementation under test (TEST INFRASTRUCTURE ONLY).

An "api" is any batch object with solve(), get(), dims(), update_bounds(l,u),
warm_start(x,y) and update_horizon(N): rldl_b200.Batch (the CUDA path through the C
ABI), oracle.ref_py.RefBatch (the unmodified reference) and
oracle.rldl_oracle.OracleBatch (the numpy restatement)."""
import numpy as np


def shift_one_stage(x, y, n, m, a, b):
    """Receding-horizon warm start: previous (x, y) moved one stage towards the origin."""
    xs, ys = np.zeros_like(x), np.zeros_like(y)
    xs[:, :n - a] = x[:, a:n]; xs[:, n - a:n] = x[:, n - a:n]
    ys[:, :m - b] = y[:, b:m]; ys[:, m - b:m] = y[:, m - b:m]
    return xs, ys


def receding_horizon(api, mdl, N0, Nmax, x0, T, shrink_every=2, feedback=None, device_shift=False):
    """solve -> apply u0 -> x0 <- A x0 + B u0 -> update_bounds ->
    warm start from the shifted previous solution -> every `shrink_every`-th step shrink the
    horizon by one stage (osqp_update_recursive), last step grow back to N0.
    `feedback[t]` (batch x nu), when given, replaces the implementation's own u0 so that
    every implementation sees identical problem data at every step.  `device_shift` uses the
    api's own warm_start_shift() (rldl_batch_warm_start_shift) instead of shifting on the host."""
    nx, nu = mdl.nx, mdl.nu
    a, b = nx + nu, nx + mdl.ny
    N, xcur, out = N0, np.array(x0, float), []
    for t in range(T):
        rc = api.solve()
        assert rc == 0, rc
        r = api.get()
        n, m = api.dims()[:2]
        out.append(dict(N=N, iter=r["iter"].copy(), status=r["status"].copy(), obj=r["obj"].copy(),
                        x=r["x"][:, :n].copy(), y=r["y"][:, :m].copy(), u0=r["x"][:, :nu].copy()))
        if t + 1 == T:
            break
        u0 = r["x"][:, :nu] if feedback is None else feedback[t]
        xcur = xcur @ mdl.A.T + u0 @ mdl.B.T
        if device_shift:
            assert api.warm_start_shift() == 0
        else:
            xs, ys = shift_one_stage(r["x"], r["y"], n, m, a, b)
        Nnew = N
        if (t + 1) % shrink_every == 0:
            Nnew = N - 1
        if t + 2 == T:
            Nnew = N0
        if Nnew != N:
            assert api.update_horizon(Nnew) == 0
            N = Nnew
        q, l, u = mdl.vectors(xcur, N, Nmax)
        assert api.update_bounds(l, u) == 0
        if not device_shift:
            assert api.warm_start(xs, ys) == 0
    return out
