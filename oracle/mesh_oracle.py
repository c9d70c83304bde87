"""TEST INFRASTRUCTURE ONLY (never imported by the product path): CPU restatement of the ground-truth speed sampling
arithmetic of dataprocessing/speed_sampling_gpu.py.

  * point_obstacle_distance (:325-336): unsigned distance of points to a triangle mesh.  The reference delegates to the
    third-party CUDA package `bvh_distance_queries` (imported at :12, not vendored in the reference tree, no version
    pinned in NTFields_env.yml), which returns the squared distance to the closest triangle; the reference takes the
    square root.  The distance to a triangle soup is independent of the search structure, so the restatement is the
    definition: exact closest point on every triangle (Voronoi-region classification, Ericson, Real-Time Collision
    Detection 5.1.5), minimum over triangles, in float64.
  * speed = clip(distance, offset, margin) / margin (:415-419).

Parity status: the reference function cannot run here (bvh_distance_queries and a CUDA device are absent), so this
restatement is pinned to closed-form cases instead (tests/test_cpu_mesh_oracle.py: distance to a box mesh, a plane, a
single triangle's vertex / edge / face regions) -- "parity unpinned" against the reference binary itself.
"""
import numpy as np


def point_triangle_dist2(p, tri):
    """p [N,3], tri [M,3,3] -> squared distances [N,M] (float64)."""
    p = np.asarray(p, np.float64)[:, None, :]
    tri = np.asarray(tri, np.float64)
    a, b, c = tri[None, :, 0], tri[None, :, 1], tri[None, :, 2]
    ab, ac, ap = b - a, c - a, p - a
    d1 = (ab * ap).sum(-1)
    d2 = (ac * ap).sum(-1)
    bp = p - b
    d3 = (ab * bp).sum(-1)
    d4 = (ac * bp).sum(-1)
    cp = p - c
    d5 = (ab * cp).sum(-1)
    d6 = (ac * cp).sum(-1)
    vc = d1 * d4 - d3 * d2
    vb = d5 * d2 - d1 * d6
    va = d3 * d6 - d5 * d4
    with np.errstate(divide="ignore", invalid="ignore"):
        den = 1.0 / (va + vb + vc)
        v = vb * den                       # interior
        w = vc * den
        todo = np.ones(v.shape, bool)

        def put(mask, vv, ww):
            nonlocal v, w, todo
            m = mask & todo
            v = np.where(m, vv, v)
            w = np.where(m, ww, w)
            todo &= ~m

        put((d1 <= 0) & (d2 <= 0), 0.0, 0.0)                                   # vertex a
        put((d3 >= 0) & (d4 <= d3), 1.0, 0.0)                                  # vertex b
        put((vc <= 0) & (d1 >= 0) & (d3 <= 0), d1 / (d1 - d3), 0.0)            # edge ab
        put((d6 >= 0) & (d5 <= d6), 0.0, 1.0)                                  # vertex c
        put((vb <= 0) & (d2 >= 0) & (d6 <= 0), 0.0, d2 / (d2 - d6))            # edge ac
        wbc = (d4 - d3) / ((d4 - d3) + (d5 - d6))
        put((va <= 0) & ((d4 - d3) >= 0) & ((d5 - d6) >= 0), 1.0 - wbc, wbc)   # edge bc
    r = ap - v[..., None] * ab - w[..., None] * ac
    return (r * r).sum(-1)


def point_obstacle_distance(points, triangles, chunk=2048):
    """Unsigned distance [N] of points [N,3] to triangles [M,3,3] (speed_sampling_gpu.py:325-336)."""
    points = np.asarray(points, np.float64)
    out = np.empty(points.shape[0], np.float64)
    for lo in range(0, points.shape[0], chunk):
        d2 = point_triangle_dist2(points[lo:lo + chunk], triangles)
        out[lo:lo + chunk] = np.sqrt(np.nanmin(d2, axis=1))
    return out


def speed_from_distance(distance, offset, margin):
    """speed_sampling_gpu.py:415-419."""
    return np.clip(distance, offset, margin) / margin


def box_mesh(lo, hi):
    """12 triangles of an axis-aligned box (test fixture)."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    v = np.array([[(hi if (k >> i) & 1 else lo)[i] for i in range(3)] for k in range(8)])
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    f = []
    for q in quads:
        f += [(q[0], q[1], q[2]), (q[0], q[2], q[3])]
    return v[np.array(f)]


def box_distance(points, lo, hi):
    """Closed form: unsigned distance to the SURFACE of an axis-aligned box."""
    p = np.asarray(points, np.float64)
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    out_d = np.maximum(np.maximum(lo - p, p - hi), 0.0)
    outside = np.sqrt((out_d ** 2).sum(-1))
    inside = np.minimum(p - lo, hi - p).min(-1)
    return np.where((out_d > 0).any(-1), outside, np.maximum(inside, 0.0))
