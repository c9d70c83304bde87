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


def train_fixture_env(seed, rows, d, margin=0.1):
    """Deterministic stand-in for one environment's dataset in the reference's on-disk layout
    (models/data_multi.py:15-31): points[rows, 2d] fp32, speed[rows, 2] in [0.1, 1], B[d, 128].  Used by
    tests/golden/make_golden_train.py (which feeds it to the unmodified reference trainer) and by the replay test."""
    rs = np.random.RandomState(seed)
    pts = rs.uniform(-0.5, 0.5, size=(rows, 2 * d)).astype(np.float32)
    half = 0.2

    def speed_of(x):                                         # box obstacle at the origin, clipped distance / margin
        q = np.abs(x) - half
        dist = np.linalg.norm(np.maximum(q, 0.0), axis=1)
        return (np.clip(dist, margin / 10.0, margin) / margin).astype(np.float32)

    spd = np.stack([speed_of(pts[:, :d]), speed_of(pts[:, d:])], axis=1)
    B = (0.5 * rs.normal(size=(d, 128))).astype(np.float32)
    return pts, spd, B


def write_multi_dataset(root, envs):
    """envs: list of (points, speed, B) -> root + str(i) + '/{sampled_points,speed,B}.npy' (models/data_multi.py:17-21)."""
    for i, (pts, spd, B) in enumerate(envs):
        p = root + str(i)
        os.makedirs(p, exist_ok=True)
        np.save(os.path.join(p, "sampled_points.npy"), pts)
        np.save(os.path.join(p, "speed.npy"), spd)
        np.save(os.path.join(p, "B.npy"), B)


def write_single_dataset(root, pts, spd):
    """models/data_mlp.py:25-31 layout (incl. the occupancy file its loader insists on)."""
    os.makedirs(root, exist_ok=True)
    np.save(os.path.join(root, "sampled_points.npy"), pts)
    np.save(os.path.join(root, "speed.npy"), spd)
    occ = np.packbits(np.zeros(128 ** 3, np.uint8))
    np.savez(os.path.join(root, "voxelized_point_cloud_128res_20000points.npz"), compressed_occupancies=occ)
