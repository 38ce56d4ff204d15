"""Full-size fixture of BASELINE config C5: status, iteration count, objective and residuals of ALL 65 536
instances of the bench workload from the UNMODIFIED reference (oracle/_ref; osqp_setup_recursive +
osqp_solve_recursive per instance, Nmax = N+1 as in make_golden.py), plus x and y of every 64th instance.

    python tests/golden/make_c5_full.py          # ~5 min on 8 cores

tests/test_fullsize_gpu.py compares the GPU results of the whole batch with it (mismatch counts)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "osqp-recursive-ldl_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

from common import ref_run  # noqa: E402
from rldl_b200 import mpc  # noqa: E402

cfg = mpc.CONFIGS["C5"]
mdl, N, total = mpc.make_model(cfg), cfg["N"], cfg["batch"]
x0 = mpc.sample_x0(mdl, total, 1234)
chunk, keep = 4096, 64
out = {k: [] for k in ("status", "iter", "obj", "pri_res", "dua_res", "x", "y")}
for lo in range(0, total, chunk):
    q, l, u = mdl.vectors(x0[lo:lo + chunk], N, N + 1)
    res = ref_run(mdl, N, N + 1, q, l, u, [("solve",), ("get",), ("dims",)])
    r, (n, m) = res[1], res[2]
    for k in ("status", "iter", "obj", "pri_res", "dua_res"):
        out[k].append(r[k])
    out["x"].append(r["x"][::keep, :n]); out["y"].append(r["y"][::keep, :m])
    print(lo, "mean iter", r["iter"].mean(), flush=True)
np.savez_compressed(os.path.join(HERE, "c5_full.npz"), batch=total, seed=1234, keep=keep, n=n, m=m,
                    status=np.concatenate(out["status"]).astype(np.int8),
                    iter=np.concatenate(out["iter"]).astype(np.int16), obj=np.concatenate(out["obj"]),
                    pri_res=np.concatenate(out["pri_res"]).astype(np.float32),
                    dua_res=np.concatenate(out["dua_res"]).astype(np.float32),
                    x=np.concatenate(out["x"]), y=np.concatenate(out["y"]))
print("wrote c5_full.npz")
