import os, sys, json
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/osqp-recursive-ldl_b200"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
from common import gpu_settings
cfg = mpc.CONFIGS["C1"]; mdl = mpc.make_model(cfg); N = cfg["N"]
for batch in (1024, 4096, 16384):
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, batch, 1234), N)
    for lat in ("0", "148", "296", "1024", "4096"):
        gb = R.Batch(mdl, N, N, q, l, u, gpu_settings())
        gb.set_tuning(lat_below=int(lat))
        ts = []
        for _ in range(6):
            gb.refactor(); gb.solve(); ts.append(gb.last_times())
        it = gb.get()["iter"]; gb.free()
        f = min(t[0] for t in ts); s = min(t[1] for t in ts)
        print(f"batch {batch} lat_below {lat}: factor {f:.3f} ms solve {s:.3f} ms  mean iter {it.mean():.1f} max {it.max()}", flush=True)
