"""Dataset readers compatible with the reference's on-disk layout.

models/data_mlp.py:25-43   : DataPath/sampled_points.npy [Ns,2d], speed.npy [Ns,2]  -> one [Ns,2d+2] tensor
models/data_multi.py:7-33  : DataPath + str(i) + /sampled_points.npy (rounded through fp16, :19), speed.npy, B.npy [d,128]
"""
import os

import numpy as np
import torch


def load_single(path):
    pts = np.load(os.path.join(path, "sampled_points.npy")).astype(np.float32)
    spd = np.load(os.path.join(path, "speed.npy")).astype(np.float32)
    return torch.from_numpy(np.concatenate([pts, spd], axis=1))


def load_multi(path, length):
    """Environment i lives in ``path + str(i)`` for i in range(length), by plain string concatenation as in
    models/data_multi.py:17 (so ``path`` is normally ``.../`` or a prefix such as ``.../env``)."""
    out = []
    for i in range(int(length)):
        p = path + str(i)
        pts = np.load(os.path.join(p, "sampled_points.npy")).astype(np.float16).astype(np.float32)
        spd = np.load(os.path.join(p, "speed.npy")).astype(np.float32)
        B = np.load(os.path.join(p, "B.npy")).astype(np.float32)
        out.append((torch.from_numpy(np.concatenate([pts, spd], axis=1)), torch.from_numpy(B)))
    return out
