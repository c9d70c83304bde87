"""Synthetic workload generator (datasets and checkpoints are unavailable offline).

Restates the sampling distribution and ground-truth speed rule of
dataprocessing/speed_sampling_gpu.py:346-365,415-419,467-475 with analytic obstacles:
  start  P ~ U[-0.5,0.5]^d ; goal = P + unit(U[-0.5,0.5]^d) * U(0, sqrt(d)), rejected outside the box;
  speed  = clip(sdf, offset, margin) / margin  at both endpoints, offset = margin/10;
  coords rounded through fp16 (models/data_multi.py:19) when fp16_round.
Obstacles: a union of axis-aligned boxes and spheres with analytic unsigned distance.
Everything is generated with torch on the requested device from a seed.
"""
import math

import torch


def default_obstacles(d, device):
    """A few boxes / spheres in the unit box (stand-in for a Gibson room / arm obstacles)."""
    g = torch.Generator().manual_seed(1234 + d)
    centers_b = (torch.rand(3, d, generator=g) - 0.5) * 0.7
    half = 0.05 + 0.1 * torch.rand(3, d, generator=g)
    centers_s = (torch.rand(3, d, generator=g) - 0.5) * 0.7
    radii = 0.05 + 0.1 * torch.rand(3, generator=g)
    return (centers_b.to(device), half.to(device), centers_s.to(device), radii.to(device))


def unsigned_distance(x, obstacles):
    cb, hb, cs, rs = obstacles
    q = (x.unsqueeze(1) - cb.unsqueeze(0)).abs() - hb.unsqueeze(0)          # [n, nb, d]
    db = torch.linalg.norm(q.clamp(min=0), dim=-1)                          # outside distance (0 inside)
    ds = (torch.linalg.norm(x.unsqueeze(1) - cs.unsqueeze(0), dim=-1) - rs.unsqueeze(0)).clamp(min=0)
    return torch.minimum(db.min(dim=1).values, ds.min(dim=1).values)


def sample_pairs(n, d, device, seed=2, margin=0.5 / 12, fp16_round=True, near_obstacle=False, obstacles=None):
    """coords[n,2d], speed[n,2] on `device`."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    if obstacles is None:
        obstacles = default_obstacles(d, device)
    offset = margin / 10.0
    out_c, out_s, have = [], [], 0
    while have < n:
        m = max(int((n - have) * 2.5), 1024)
        P = torch.rand(m, d, device=device, generator=gen) - 0.5
        dirs = torch.rand(m, d, device=device, generator=gen) - 0.5
        dirs = dirs / torch.linalg.norm(dirs, dim=1, keepdim=True)
        length = torch.rand(m, 1, device=device, generator=gen) * math.sqrt(d)
        G = P + dirs * length
        keep = (G.abs() <= 0.5).all(dim=1)
        if near_obstacle:
            dist0 = unsigned_distance(P, obstacles)
            keep &= (dist0 > offset) & (dist0 < margin)
        P, G = P[keep], G[keep]
        c = torch.cat([P, G], dim=1)
        if fp16_round:
            c = c.to(torch.float16).to(torch.float32)
        s0 = unsigned_distance(c[:, :d], obstacles).clamp(offset, margin) / margin
        s1 = unsigned_distance(c[:, d:], obstacles).clamp(offset, margin) / margin
        out_c.append(c)
        out_s.append(torch.stack([s0, s1], dim=1))
        have += c.shape[0]
    return torch.cat(out_c)[:n].contiguous(), torch.cat(out_s)[:n].contiguous()


def sample_B(d, device, seed=1, E=1):
    """B = 0.5*N(0,1), [E,d,128] (dataprocessing/speed_sampling_gpu.py:493)."""
    g = torch.Generator().manual_seed(seed)
    return (0.5 * torch.randn(E, d, 128, generator=g)).to(device)


def sample_queries(Q, d, device, seed=3):
    """start/goal ~ U[-0.4,0.4]^d."""
    g = torch.Generator().manual_seed(seed)
    return ((torch.rand(Q, 2 * d, generator=g) - 0.5) * 0.8).to(device)


BOX_SCENE = (((-0.3, -0.3, -0.3), (-0.05, 0.0, 0.1)), ((0.1, -0.2, -0.4), (0.4, 0.35, -0.1)),
             ((-0.1, 0.2, 0.15), (0.3, 0.45, 0.4)))


def box_scene_mesh(device, levels=6, boxes=BOX_SCENE, shuffle_seed=11):
    """Synthetic obstacle mesh for the speed-sampling benchmark: the surfaces of a few axis-aligned boxes, every
    triangle split into four `levels` times (36 * 4**levels triangles), in shuffled order as a mesh file would be.
    Returns triangles [M,3,3] fp32 on `device`."""
    quads = ((0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3))
    tris = []
    for lo, hi in boxes:
        v = torch.tensor([[(hi if (k >> i) & 1 else lo)[i] for i in range(3)] for k in range(8)], dtype=torch.float64)
        for q in quads:
            tris.append(v[[q[0], q[1], q[2]]])
            tris.append(v[[q[0], q[2], q[3]]])
    t = torch.stack(tris).to(device)
    for _ in range(levels):
        a, b, c = t[:, 0], t[:, 1], t[:, 2]
        ab, bc, ca = (a + b) / 2, (b + c) / 2, (c + a) / 2
        t = torch.cat([torch.stack(x, 1) for x in ((a, ab, ca), (ab, b, bc), (ca, bc, c), (ab, bc, ca))])
    g = torch.Generator(device=device).manual_seed(shuffle_seed)
    return t[torch.randperm(t.shape[0], device=device, generator=g)].to(torch.float32).contiguous()
