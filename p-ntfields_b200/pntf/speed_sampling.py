"""Ground-truth speed sampling for point robots: the data-preparation step that writes the
`sampled_points.npy` / `speed.npy` / `B.npy` files the trainers read.

Mirrors dataprocessing/speed_sampling_gpu.py of the reference (same function names, argument meaning and outputs):
  point_obstacle_distance            :325-336   unsigned distance of points to the obstacle mesh
  point_append_list                  :338-392   rejection sampler for (x0, x1) pairs near the obstacle boundary
  point_rand_sample_bound_points     :395-421   wrapper: numpy mesh in, (sampled_points, speed) numpy out
  sample_speed                       :423-495   reads <name>_scaled.off, samples, saves the three .npy files

The distance queries run in libpntf_b200.so (`pntf_mesh_distance`, hand-written CUDA) instead of the reference's
third-party `bvh_distance_queries`; random numbers come from torch's CUDA generator exactly as in the reference, so
a seeded run draws the same candidate stream.  The arm branch (:29-323) lives in pntf/arm_sampling.py (own serial-chain
forward kinematics instead of pytorch_kinematics; the UR5's URDF / link meshes are not available offline, so
`sample_speed` cannot read them and asks the caller for a chain and link meshes).  There is no CPU fallback.
"""
import os

import numpy as np
import torch

from . import capi


def _morton_order(p):
    """Permutation that sorts points [N,3] along a 30-bit Morton curve of their bounding box (host-side plumbing:
    spatially coherent CTAs skip most triangle tiles; results do not depend on the order)."""
    lo = p.min(dim=0).values
    span = (p.max(dim=0).values - lo).clamp_min(1e-20)
    q = ((p - lo) / span * 1023.0).to(torch.int64).clamp_(0, 1023)

    def spread(v):
        v = (v | (v << 16)) & 0x030000FF
        v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3
        v = (v | (v << 2)) & 0x09249249
        return v

    code = spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
    return torch.argsort(code)


class ObstacleMesh:
    """Triangle soup [M,3,3] on the GPU, sorted once along a Morton curve of the triangle centroids."""

    def __init__(self, triangles_obs):
        t = triangles_obs.reshape(-1, 3, 3).to(torch.float32)
        if not t.is_cuda:
            raise RuntimeError("the obstacle mesh must be on a CUDA device (no CPU fallback)")
        self.triangles = t[_morton_order(t.mean(dim=1))].contiguous()

    def distance(self, query_points, offset=None, margin=None):
        p = query_points.to(torch.float32)
        order = _morton_order(p)
        res = capi.mesh_distance(self.triangles, p[order].contiguous(), offset, margin)
        inv = torch.empty_like(order)
        inv[order] = torch.arange(order.numel(), device=order.device)
        if isinstance(res, tuple):
            return res[0][inv], res[1][inv]
        return res[inv]


def point_obstacle_distance(query_points, triangles_obs):
    """query_points [N,3], triangles_obs [1,M,3,3] (or an ObstacleMesh) -> unsigned distance [N] (:325-336)."""
    mesh = triangles_obs if isinstance(triangles_obs, ObstacleMesh) else ObstacleMesh(triangles_obs)
    return mesh.distance(query_points)


def point_append_list(X_list, Y_list, triangles_obs, numsamples, dim, offset, margin):
    """Rejection sampler (:338-392): draw 8*numsamples start points P in the unit box and partners nP = P + r * u
    (u a random direction, r ~ U[0, sqrt(dim)]); keep pairs whose partner is inside the box and whose start point lies
    in the band offset < distance < margin around the obstacles; append [x0 x1] and [d0 d1] until enough were found."""
    mesh = triangles_obs if isinstance(triangles_obs, ObstacleMesh) else ObstacleMesh(triangles_obs)
    dev = mesh.triangles.device
    OutsideSize = numsamples + 2
    WholeSize = 0
    while OutsideSize > 0:
        P = torch.rand((8 * numsamples, dim), dtype=torch.float32, device=dev) - 0.5
        dP = torch.rand((8 * numsamples, dim), dtype=torch.float32, device=dev) - 0.5
        rL = torch.rand((8 * numsamples, 1), dtype=torch.float32, device=dev) * np.sqrt(dim)
        nP = P + torch.nn.functional.normalize(dP, dim=1) * rL
        PointsInside = torch.all((nP <= 0.5), dim=1) & torch.all((nP >= -0.5), dim=1)
        x0 = P[PointsInside, :]
        x1 = nP[PointsInside, :]
        if x0.shape[0] <= 1:
            continue
        obs_distance0 = mesh.distance(x0)
        where_d = (obs_distance0 > offset) & (obs_distance0 < margin)
        x0 = x0[where_d]
        x1 = x1[where_d]
        y0 = obs_distance0[where_d]
        y1 = mesh.distance(x1) if x1.shape[0] else y0
        X_list.append(torch.cat((x0, x1), 1))
        Y_list.append(torch.cat((y0.unsqueeze(1), y1.unsqueeze(1)), 1))
        OutsideSize = OutsideSize - x0.shape[0]
        WholeSize = WholeSize + x0.shape[0]
        if WholeSize > numsamples:
            break
    return X_list, Y_list


def point_rand_sample_bound_points(numsamples, dim, v_obs, f_obs, offset, margin, device="cuda"):
    """numpy vertices / faces in, (sampled_points [n, 2*dim], speed [n, 2]) numpy out (:395-421)."""
    numsamples = int(numsamples)
    v_obs = torch.as_tensor(np.asarray(v_obs), dtype=torch.float32, device=device)
    f_obs = torch.as_tensor(np.asarray(f_obs), dtype=torch.long, device=device)
    mesh = ObstacleMesh(v_obs[f_obs].unsqueeze(dim=0))
    X_list, Y_list = point_append_list([], [], mesh, numsamples, dim, offset, margin)
    X = torch.cat(X_list, 0)[:numsamples]
    Y = torch.cat(Y_list, 0)[:numsamples]
    sampled_points = X.detach().cpu().numpy()
    distance = Y.detach().cpu().numpy()
    speed = np.zeros((distance.shape[0], 2))
    speed[:, 0] = np.clip(distance[:, 0], a_min=offset, a_max=margin) / margin
    speed[:, 1] = np.clip(distance[:, 1], a_min=offset, a_max=margin) / margin
    return sampled_points, speed


def read_off(path):
    """Minimal OFF reader (the reference uses igl.read_triangle_mesh, :472): vertices [V,3] float64, faces [F,3] int64
    (polygons are fan-triangulated)."""
    with open(path) as f:
        tok = f.read().split()
    if not tok or not tok[0].startswith("OFF"):
        raise ValueError("%s: not an OFF file" % path)
    pos = 1
    if tok[0] != "OFF":                      # "OFF8 6 0" on one line
        tok = [tok[0][:3], tok[0][3:]] + tok[1:]
    nv, nf = int(tok[pos]), int(tok[pos + 1])
    pos += 3
    v = np.array(tok[pos:pos + 3 * nv], dtype=np.float64).reshape(nv, 3)
    pos += 3 * nv
    faces = []
    for _ in range(nf):
        k = int(tok[pos])
        idx = [int(t) for t in tok[pos + 1:pos + 1 + k]]
        pos += 1 + k
        for j in range(1, k - 1):
            faces.append((idx[0], idx[j], idx[j + 1]))
    return v, np.array(faces, dtype=np.int64).reshape(-1, 3)


def band_for_task(task_name, limit=0.5):
    """(offset, margin) of the sampling band per dataset family (:457-466)."""
    margin = limit / 5.0 if task_name in ("c3d", "test") else limit / 12.0
    return margin / 10.0, margin


def sample_speed(path, numsamples, dim, task_name=None, device="cuda"):
    """Sample `numsamples` pairs around the mesh `<dir>/<name>_scaled.off` next to `path` and save
    sampled_points.npy, speed.npy and B.npy into that directory (:423-495).  task_name defaults to the third path
    component as in the reference (`./datasets/<task>/...`)."""
    out_path = os.path.dirname(path)
    if task_name is None:
        parts = out_path.split("/")
        task_name = parts[2] if len(parts) > 2 else ""
    if task_name == "arm":
        raise NotImplementedError("arm datasets: call pntf.arm_sampling.arm_rand_sample_bound_points with a SerialChain "
                                  "(SerialChain.from_urdf) and the links' collision-mesh vertices; the UR5 files the "
                                  "reference reads from ./datasets/arm/UR5 are not part of this repository")
    file_name = os.path.splitext(os.path.basename(path))[0]
    input_file = os.path.join(out_path, file_name + "_scaled.off")
    offset, margin = band_for_task(task_name)
    v, f = read_off(input_file)
    sampled_points, speed = point_rand_sample_bound_points(numsamples, dim, v, f, offset, margin, device=device)
    B = 0.5 * np.random.normal(0, 1, size=(3, 128))
    np.save("{}/sampled_points".format(out_path), sampled_points)
    np.save("{}/speed".format(out_path), speed)
    np.save("{}/B".format(out_path), B)
    return sampled_points, speed, B
