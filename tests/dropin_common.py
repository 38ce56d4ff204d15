"""Helpers for the drop-in tests (test infrastructure): run the user program of
tests/dropin/mpc_user.c and parse what it prints."""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden", "dropin_mpc_user_ref.txt")
GOLDEN_API = os.path.join(ROOT, "tests", "golden", "dropin_api_user_ref.txt")


def binary(name):
    return os.path.join(BIN, name)


def run(name, timeout=300):
    return subprocess.run([binary(name)], capture_output=True, text=True, timeout=timeout)


def parse(text):
    """-> {"rc": {line: value}, "dims": (n, m), "solves": {tag: {n, m, status_val, iter, obj, ..., status, x, y}}}"""
    out = {"solves": {}, "rc": {}}
    for line in text.splitlines():
        w = line.split()
        if not w:
            continue
        if w[0] in ("SETUP", "BADBOUNDS", "BADRHO", "CLEANUP", "DONE") and "rc" in w:
            out["rc"][" ".join(w[:w.index("rc")])] = int(w[w.index("rc") + 1])
        elif w[0] == "DIMS":
            out["dims"] = (int(w[2]), int(w[4]))
        elif w[0] == "SOLVE":
            rec = out["solves"].setdefault(w[1], {})
            k = 2
            while k + 1 < len(w) and w[k] != "status":
                rec[w[k]] = float(w[k + 1]) if w[k] in ("obj", "pri", "dua") else int(w[k + 1])
                k += 2
            rec["status"] = line.split('"')[1]
        elif w[0] == "POLISH":
            out["solves"].setdefault(w[1], {})["status_polish"] = int(w[3])
        elif w[0] in ("X", "Y") and len(w) > 1:
            out["solves"].setdefault(w[1], {})[w[0].lower()] = np.array([float(v) for v in w[2:]])
    return out


def rel(a, b):
    if np.asarray(b).size == 0:
        return 0.0 if np.asarray(a).size == 0 else np.inf
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-12))


def assert_same_run(got, want, tol, what):
    """Same return codes, dimensions, status and iteration count at every solve; x, y, obj within tol."""
    assert got["rc"] == want["rc"], (what, got["rc"], want["rc"])
    assert got.get("dims") == want.get("dims"), (what, got.get("dims"), want.get("dims"))
    assert list(got["solves"]) == list(want["solves"]) and want["solves"], (what, list(got["solves"]))
    for tag, w in want["solves"].items():
        g = got["solves"][tag]
        for key in ("rc", "n", "m", "status_val", "iter", "status", "status_polish"):
            assert g.get(key) == w.get(key), (what, tag, key, g.get(key), w.get(key))
        assert g["x"].shape == (w["n"],) and g["y"].shape == (w["m"],), (what, tag)
        assert rel(g["x"], w["x"]) < tol, (what, tag, "x", rel(g["x"], w["x"]))
        assert rel(g["y"], w["y"]) < tol, (what, tag, "y", rel(g["y"], w["y"]))
        assert abs(g["obj"] - w["obj"]) <= tol * max(abs(w["obj"]), 1e-12), (what, tag, "obj", g["obj"], w["obj"])
