"""Helpers for the drop-in tests (test infrastructure): run the user program of
tests/dropin/mpc_user.c and parse what it prints."""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden", "dropin_mpc_user_ref.txt")
TAGS = ["cold", "step1", "step2", "lincost", "shrink", "grow", "tight"]


def binary(name):
    return os.path.join(BIN, name)


def run(name, timeout=300):
    return subprocess.run([binary(name)], capture_output=True, text=True, timeout=timeout)


def parse(text):
    """-> {"setup_rc", "dims", "badbounds_rc", "done_rc", "solves": {tag: {...}}}"""
    out = {"solves": {}}
    for line in text.splitlines():
        w = line.split()
        if not w:
            continue
        if w[0] == "SETUP":
            out["setup_rc"] = int(w[2])
        elif w[0] == "DIMS":
            out["dims"] = (int(w[2]), int(w[4]))
        elif w[0] == "BADBOUNDS":
            out["badbounds_rc"] = int(w[2])
        elif w[0] == "DONE":
            out["done_rc"] = int(w[2])
        elif w[0] == "SOLVE":
            rec = out["solves"].setdefault(w[1], {})
            rec.update(n=int(w[3]), m=int(w[5]), status_val=int(w[7]), iter=int(w[9]), obj=float(w[11]),
                       pri=float(w[13]), dua=float(w[15]), status=line.split('"')[1])
        elif w[0] in ("X", "Y") and len(w) > 1 and w[1] in TAGS:
            out["solves"].setdefault(w[1], {})[w[0].lower()] = np.array([float(v) for v in w[2:]])
    return out


def rel(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-12))


def assert_same_run(got, want, tol, what):
    """Same return codes, dimensions, status and iteration count at every solve; x, y, obj within tol."""
    for key in ("setup_rc", "dims", "badbounds_rc", "done_rc"):
        assert got.get(key) == want.get(key), (what, key, got.get(key), want.get(key))
    assert list(got["solves"]) == TAGS == list(want["solves"]), (what, list(got["solves"]))
    for tag in TAGS:
        g, w = got["solves"][tag], want["solves"][tag]
        for key in ("n", "m", "status_val", "iter", "status"):
            assert g[key] == w[key], (what, tag, key, g[key], w[key])
        assert g["x"].shape == (w["n"],) and g["y"].shape == (w["m"],), (what, tag)
        assert rel(g["x"], w["x"]) < tol, (what, tag, "x", rel(g["x"], w["x"]))
        assert rel(g["y"], w["y"]) < tol, (what, tag, "y", rel(g["y"], w["y"]))
        assert abs(g["obj"] - w["obj"]) <= tol * max(abs(w["obj"]), 1e-12), (what, tag, "obj", g["obj"], w["obj"])
