"""Dataset readers compatible with the reference's on-disk layout.

models/data_mlp.py:25-43   : DataPath/sampled_points.npy [Ns,2d], speed.npy [Ns,2]  -> one [Ns,2d+2] tensor
models/data_multi.py:7-33  : DataPath/<env>/sampled_points.npy (rounded through fp16, :19), speed.npy, B.npy [d,128]
"""
import os

import numpy as np
import torch


def load_single(path):
    pts = np.load(os.path.join(path, "sampled_points.npy")).astype(np.float32)
    spd = np.load(os.path.join(path, "speed.npy")).astype(np.float32)
    return torch.from_numpy(np.concatenate([pts, spd], axis=1))


def load_multi(path, length):
    envs = sorted(d for d in os.listdir(path) if os.path.isdir(os.path.join(path, d)))[:length]
    out = []
    for e in envs:
        p = os.path.join(path, e)
        pts = np.load(os.path.join(p, "sampled_points.npy")).astype(np.float16).astype(np.float32)
        spd = np.load(os.path.join(p, "speed.npy")).astype(np.float32)
        B = np.load(os.path.join(p, "B.npy")).astype(np.float32)
        out.append((torch.from_numpy(np.concatenate([pts, spd], axis=1)), torch.from_numpy(B)))
    return out
