"""Experiment: iteration-count distribution of the C5 workload and what instance ordering is worth."""
import os, sys, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "osqp-recursive-ldl_b200"))
import numpy as np
import rldl_b200 as R
from rldl_b200 import mpc
import bench

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
cfg, mdl = bench.workload("C5", batch)
N = cfg["N"]
x0 = mpc.sample_x0(mdl, batch, bench.SEED)
q, l, u = mdl.vectors(x0, N)
st = R.default_settings(**bench.SETTINGS)

def run(order, tag):
    gb = R.Batch(mdl, N, N, q[order], l[order], u[order], st)
    ts = []
    for _ in range(3):
        gb.refactor(); gb.solve(); ts.append(gb.last_times()[1])
    out = gb.get()
    gb.free()
    print(tag, "solve ms", [round(t, 1) for t in ts], flush=True)
    return out["iter"], min(ts)

it, t_nat = run(np.arange(batch), "natural")
pc = [50, 75, 90, 95, 99, 99.9, 100]
print("iters mean", it.mean(), "percentiles", dict(zip(pc, np.percentile(it, pc).tolist())))
print("sum iters", it.sum(), " frac of work in instances above p90:", it[it > np.percentile(it, 90)].sum() / it.sum())
desc = np.argsort(-it, kind="stable")
_, t_desc = run(desc, "longest-first")
_, t_asc = run(desc[::-1], "shortest-first")
# correlation of iteration count with simple features
nx0 = np.abs(x0).max(axis=1)
print("corr(iter, |x0|inf)", np.corrcoef(it, nx0)[0, 1], "corr(iter,|x0|2)", np.corrcoef(it, np.linalg.norm(x0, axis=1))[0, 1])
np.save(os.path.join(ROOT, "gpurun_out", "iters_c5.npy"), it.astype(np.int32))
print(json.dumps({"natural_ms": t_nat, "longest_first_ms": t_desc, "shortest_first_ms": t_asc}))
