"""CPU suite: the C-ABI library loads and exports every symbol include/osqp_rldl_abi.h
declares; struct layouts match the reference's; entry points fail loudly without a device;
the multi-GPU sharding logic (world_size 2, gloo)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import rldl_b200 as R
from oracle import ref_py

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "osqp_rldl_abi.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?(?:c_int|void|char|csc|RLDLBatch)\s*\*?\s*(\w+)\s*\(", txt, flags=re.M)
    return sorted(set(n for n in names if not n.startswith("(")))


def test_library_exports_every_declared_symbol():
    lib = R.lib()
    syms = header_symbols()
    assert len(syms) >= 60, syms
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    assert sorted(set(R.ABI_SYMBOLS)) == syms, set(R.ABI_SYMBOLS) ^ set(syms)
    assert b"recursive-LDL" in lib.rldl_version()


def test_default_settings_match_reference_constants():
    s = R.default_settings()   # reference include/constants.h:59-114
    assert (s.rho, s.sigma, s.scaling, s.alpha, s.max_iter) == (0.1, 1e-6, 10, 1.6, 4000)
    assert (s.eps_abs, s.eps_rel, s.eps_prim_inf, s.eps_dual_inf) == (1e-3, 1e-3, 1e-4, 1e-4)
    assert (s.adaptive_rho, s.adaptive_rho_interval, s.adaptive_rho_tolerance, s.adaptive_rho_fraction) == (1, 0, 5.0, 0.4)
    assert (s.check_termination, s.warm_start, s.polish, s.verbose) == (25, 1, 0, 1)


@pytest.mark.skipif(not ref_py.available(), reason="oracle/_ref not built")
def test_settings_struct_layout_matches_reference_build():
    assert C.sizeof(R.OSQPSettings) == ref_py.lib().refb_sizeof_settings()
    a, b = R.default_settings(), ref_py.default_settings()
    for name, _ in R.OSQPSettings._fields_:
        assert getattr(a, name) == getattr(b, name), name


def test_setup_fails_loudly_without_a_device():
    lib = R.lib()
    if lib.rldl_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from rldl_b200 import mpc
    mdl = mpc.make_model(mpc.CONFIGS["C1"])
    q, l, u = mdl.vectors(mpc.sample_x0(mdl, 2), 20)
    with pytest.raises(RuntimeError):   # OSQP_LINSYS_SOLVER_LOAD_ERROR: no CPU path exists
        R.Batch(mdl, 20, 20, q, l, u, R.default_settings(scaling=0, adaptive_rho=0))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "osqp-recursive-ldl_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.replace("oracle/_ref", "").lower() or f == "mpc.py", os.path.join(dp, f)


_WORKER = r"""
import os, sys, json
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
import bench
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
lo, hi = bench.shard_range(rank, world, 65536 + 3)
t = torch.tensor([float(hi - lo)], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.SUM)
mx = torch.tensor([1.0 + rank], dtype=torch.float64)
dist.all_reduce(mx, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps(dict(total=t.item(), mx=mx.item(), lo=lo, hi=hi)))
dist.destroy_process_group()
"""


def test_sharding_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER.format(root=ROOT))
    cp = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                         "--master-addr", "127.0.0.1", "--master-port", "29571", str(script)],
                        capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert cp.returncode == 0, cp.stderr[-2000:]
    import json
    out = json.loads([ln for ln in cp.stdout.splitlines() if ln.startswith("{")][-1])
    assert out["total"] == 65536 + 3 and out["mx"] == 2.0 and out["lo"] == 0


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    cp = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                        capture_output=True, text=True, env=env, timeout=120)
    assert cp.returncode == 0 and cp.stdout.strip() == ""
