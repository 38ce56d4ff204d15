"""The -DRLDL_DEBUG build of the library (make -C osqp-recursive-ldl_b200 debug): the same sources with in-kernel
bounds asserts on ring slots, stage-record offsets, scheduler slots and instance ids (compute-sanitizer is closed on
this pool).  tests/debug_paths.py takes the hot paths through it in a fresh process — streaming kernel with
pre-emption, concurrent lane, tail hand-off, latency-mode kernel, horizon shrink / grow, adaptive-rho park / resume —
and must end without a trap and with the digests of the release build."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEBUG_LIB = os.path.join(ROOT, "osqp-recursive-ldl_b200", "lib", "librldl_b200_debug.so")


def _run(env_lib):
    env = dict(os.environ)
    if env_lib:
        env["RLDL_B200_LIB"] = env_lib
    else:
        env.pop("RLDL_B200_LIB", None)
    cp = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "debug_paths.py")], capture_output=True, text=True,
                        timeout=400, env=env)
    assert cp.returncode == 0 and "PATHS DONE" in cp.stdout, (cp.returncode, cp.stdout[-1500:], cp.stderr[-1500:])
    assert "RLDL_ASSERT" not in cp.stdout + cp.stderr
    return cp.stdout


def test_hot_paths_run_clean_on_the_assert_build_and_agree_with_the_release_build():
    assert os.path.exists(DEBUG_LIB), "debug library not built: make -C osqp-recursive-ldl_b200 debug"
    dbg, rel = _run(DEBUG_LIB), _run(None)
    assert "LIB librldl_b200_debug.so" in dbg and "LIB librldl_b200.so" in rel
    d = {ln.split()[1]: ln.split()[2:] for ln in dbg.splitlines() if ln.startswith("DIGEST")}
    r = {ln.split()[1]: ln.split()[2:] for ln in rel.splitlines() if ln.startswith("DIGEST")}
    assert list(d) == list(r) and len(d) == 7
    for tag in d:
        # status histogram and iteration sum identical; |x|, |y| sums to 1e-9 relative (which kernel finishes an
        # instance is timing dependent: rounding-level differences)
        assert d[tag][:d[tag].index("x")] == r[tag][:r[tag].index("x")], (tag, d[tag], r[tag])
        for key in ("x", "y"):
            a, b = float(d[tag][d[tag].index(key) + 1]), float(r[tag][r[tag].index(key) + 1])
            assert abs(a - b) <= 1e-9 * max(abs(b), 1.0), (tag, key, a, b)
