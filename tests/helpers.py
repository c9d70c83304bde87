"""Shared helpers for the test-suite (fixture loading, digest comparison)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: z[k] for k in z.files}


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


def check_digest(fx, prefix, name, t, rtol):
    """Compare a parameter-shaped tensor with the digest tests/golden/make_golden.py stored."""
    t = np.asarray(t, np.float32)
    key = prefix + name
    if t.ndim == 1:
        assert rel_err(t, fx[key]) < rtol, (key, rel_err(t, fx[key]))
        return
    rows = fx[key + "#rowidx"]
    scale = np.abs(fx[key + "#rows"]).max() + 1e-30
    assert np.abs(t[rows] - fx[key + "#rows"]).max() / scale < rtol, key
    assert rel_err(t.sum(1), fx[key + "#rowsum"]) < 4 * rtol, key
    assert rel_err(t.sum(0), fx[key + "#colsum"]) < 4 * rtol, key
    fro = np.sqrt((t.astype(np.float64) ** 2).sum())
    assert abs(fro - fx[key + "#fro"]) / fx[key + "#fro"] < rtol, key


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))
