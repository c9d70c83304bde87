"""Ground-truth speed sampling for an articulated arm (arm branch): mirrors
dataprocessing/speed_sampling_gpu.py:29-323 of the reference.

  bbox_corner / separate_axis        :29-70     oriented-box corners and the 15 separating-axis candidates
  arm_obstacle_obb                   :76-151    OBB pre-filter of the arm's links against the obstacle's box
  arm_obstacle_distance              :153-218   per configuration: forward kinematics, every vertex of every link's
                                                collision mesh transformed into the world, unsigned distance of all of
                                                them to the obstacle mesh, minimum over the arm
  arm_append_list                    :220-287   rejection sampler in joint space (configurations = scale * x, scale = pi / 0.5)
  arm_rand_sample_bound_points       :289-323   wrapper: numpy obstacle mesh in, (sampled_points, speed) numpy out

The reference builds its kinematic chain with the third-party `pytorch_kinematics` from a URDF and reads the UR5's
collision meshes with `igl`; neither the packages nor the robot files are available here.  `SerialChain` is a
from-scratch batched serial-chain forward kinematics (revolute / prismatic / fixed joints, origin xyz + rpy, axis) with a
small URDF reader (`SerialChain.from_urdf`), and link meshes are plain vertex arrays (`box_link_vertices` builds
box-shaped stand-ins).  The distance queries are the same `pntf_mesh_distance` kernel the point-robot branch uses (the
reference calls `bvh_distance_queries` in both).  Differences that are deliberate: the obstacle box of the OBB filter
is an argument (the reference hard-codes one, :103) and link boxes are passed in (the reference reads `<link>bbox.npy`).
There is no CPU fallback.
"""
import math
import xml.etree.ElementTree as ET

import numpy as np
import torch

from .speed_sampling import ObstacleMesh


def _rpy_matrix(rpy):
    r, p, y = [float(v) for v in rpy]
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]], np.float64)


class SerialChain:
    """Batched forward kinematics of a serial chain.  joints: list of dicts
    {name (child link), type ('revolute' | 'continuous' | 'prismatic' | 'fixed'), xyz, rpy, axis}."""

    def __init__(self, joints, base_link="base_link", device="cuda", dtype=torch.float32):
        self.joints = [dict(j) for j in joints]
        self.base_link = base_link
        self.device, self.dtype = torch.device(device), dtype
        self.origins = []
        for j in self.joints:
            T = np.eye(4)
            T[:3, :3] = _rpy_matrix(j.get("rpy", (0, 0, 0)))
            T[:3, 3] = j.get("xyz", (0, 0, 0))
            self.origins.append(torch.tensor(T, dtype=dtype, device=self.device))
        self.n_dof = sum(j["type"] != "fixed" for j in self.joints)

    @classmethod
    def from_urdf(cls, text, end_link, device="cuda", dtype=torch.float32):
        """The chain root -> end_link of a URDF (what pk.build_serial_chain_from_urdf(text, end_link) selects, :298-300)."""
        root = ET.fromstring(text)
        by_child = {}
        for j in root.findall("joint"):
            o = j.find("origin")
            ax = j.find("axis")
            by_child[j.find("child").get("link")] = {
                "name": j.find("child").get("link"), "parent": j.find("parent").get("link"), "type": j.get("type"),
                "xyz": [float(v) for v in (o.get("xyz", "0 0 0") if o is not None else "0 0 0").split()],
                "rpy": [float(v) for v in (o.get("rpy", "0 0 0") if o is not None else "0 0 0").split()],
                "axis": [float(v) for v in (ax.get("xyz") if ax is not None else "1 0 0").split()]}
        chain, link = [], end_link
        while link in by_child:
            chain.append(by_child[link])
            link = by_child[link]["parent"]
        chain.reverse()
        return cls(chain, base_link=link, device=device, dtype=dtype)

    def to(self, dtype=None, device=None):
        return SerialChain(self.joints, self.base_link, device or self.device, dtype or self.dtype)

    def forward_kinematics(self, th, end_only=False):
        """th [N, n_dof] -> {link name: world transform [N,4,4]} in chain order, the base link first (identity)."""
        th = th.to(device=self.device, dtype=self.dtype)
        N = th.shape[0]
        T = torch.eye(4, dtype=self.dtype, device=self.device).expand(N, 4, 4).contiguous()
        out = {self.base_link: T}
        k = 0
        for j, O in zip(self.joints, self.origins):
            T = T @ O
            if j["type"] != "fixed":
                q = th[:, k]
                k += 1
                a = torch.tensor(j["axis"], dtype=self.dtype, device=self.device)
                a = a / torch.linalg.norm(a)
                M = torch.eye(4, dtype=self.dtype, device=self.device).repeat(N, 1, 1)
                if j["type"] == "prismatic":
                    M[:, :3, 3] = q[:, None] * a[None, :]
                else:                                        # Rodrigues: R = I + sin q K + (1 - cos q) K^2
                    K = torch.zeros(3, 3, dtype=self.dtype, device=self.device)
                    K[0, 1], K[0, 2], K[1, 0], K[1, 2], K[2, 0], K[2, 1] = -a[2], a[1], a[2], -a[0], -a[1], a[0]
                    s, c = torch.sin(q)[:, None, None], torch.cos(q)[:, None, None]
                    M[:, :3, :3] = torch.eye(3, dtype=self.dtype, device=self.device) + s * K + (1 - c) * (K @ K)
                T = T @ M
            out[j["name"]] = T
        if end_only:
            return T
        return out


def box_link_vertices(lo, hi, n=3):
    """Vertices of a box-shaped link collision mesh: an n x n grid on each face (stand-in for `<link>.obj`, :170)."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    t = np.linspace(0.0, 1.0, n)
    pts = []
    for ax in range(3):
        for side in (0.0, 1.0):
            u, v = np.meshgrid(t, t)
            p = np.zeros((n * n, 3))
            p[:, ax] = side
            p[:, (ax + 1) % 3] = u.ravel()
            p[:, (ax + 2) % 3] = v.ravel()
            pts.append(lo + p * (hi - lo))
    return np.unique(np.concatenate(pts), axis=0).astype(np.float32)


def bbox_corner(bbox):
    """bbox = (max xyz, min xyz) as the reference stores it -> homogeneous corners [4, 8] (:29-43)."""
    b = torch.as_tensor(bbox, dtype=torch.float32)
    hi, lo = b[:3], b[3:]
    c = torch.ones(4, 8, dtype=torch.float32, device=b.device)
    for k in range(8):
        for i in range(3):
            c[i, k] = hi[i] if (k >> i) & 1 else lo[i]
    return c


def separate_axis(transform):
    """The 15 separating-axis candidates of a box with orientation transform[:, :3, :3] against a world-aligned box:
    3 world axes, 3 box axes (the rows, as the reference takes them, :63), 9 cross products (:46-70)."""
    R = transform[:, 0:3, 0:3]
    N = R.shape[0]
    eye = torch.eye(3, dtype=R.dtype, device=R.device)
    axis = torch.zeros(N, 15, 3, dtype=R.dtype, device=R.device)
    axis[:, 0:3] = eye
    axis[:, 3:6] = R
    for i in range(3):
        axis[:, 6 + 3 * i:9 + 3 * i] = torch.cross(R[:, i:i + 1, :].expand(N, 3, 3), eye.expand(N, 3, 3), dim=2)
    return axis


def arm_obstacle_obb(th_batch, chain, link_boxes, obstacle_box, margin, first_link=3):
    """OBB pre-filter (:76-151).  link_boxes: one (max xyz, min xyz) box per link from chain position `first_link` on
    (grown by `margin`, :97-98); obstacle_box: (max xyz, min xyz) of the obstacle.  Returns a bool [N]: no separating
    axis was found between the obstacle and the LAST link's box -- what the reference returns (`where_list.append(nonsep)`,
    :149; its running any-link flag `where_coll` is computed and dropped), kept as is."""
    dev = chain.device
    obs = bbox_corner(torch.as_tensor(obstacle_box, dtype=torch.float32)).to(dev)
    corners = []
    for b in link_boxes:
        b = torch.as_tensor(b, dtype=torch.float32).clone()
        b[:3] += margin
        b[3:] -= margin
        corners.append(bbox_corner(b).to(dev))
    tg = chain.forward_kinematics(th_batch)
    names = list(tg.keys())
    nonsep = torch.ones(th_batch.shape[0], dtype=torch.bool, device=dev)
    for it, name in enumerate(names):
        if it < first_link or it - first_link >= len(corners):
            continue
        m = tg[name]
        rob = m @ corners[it - first_link]
        axis = separate_axis(m)
        ra = axis @ rob[:, 0:3, :]
        oa = axis @ obs[0:3, :]
        sep = (ra.max(dim=2).values < oa.min(dim=2).values) | (oa.max(dim=2).values < ra.min(dim=2).values)
        nonsep = ~sep.any(dim=1)
    return nonsep


def arm_obstacle_distance(th_batch, chain, link_vertices, triangles_obs, first_link=2, batch_size=50000):
    """Smallest unsigned distance between the arm and the obstacle mesh per configuration (:153-218).
    link_vertices: {link name: vertices [V,3]} of the collision meshes; links before chain position `first_link`
    (the base and the first link, `iter > 1` at :168) are skipped as in the reference."""
    mesh = triangles_obs if isinstance(triangles_obs, ObstacleMesh) else ObstacleMesh(triangles_obs)
    dev = mesh.triangles.device
    out = []
    for lo in range(0, th_batch.shape[0], batch_size):
        th = th_batch[lo:lo + batch_size]
        tg = chain.forward_kinematics(th)
        plist = []
        for it, name in enumerate(tg):
            if it < first_link or name not in link_vertices:
                continue
            v = torch.as_tensor(link_vertices[name], dtype=torch.float32, device=dev)
            hv = torch.cat([v, torch.ones(v.shape[0], 1, device=dev)], dim=1)            # [V,4]
            p = torch.matmul(tg[name], hv.T).permute(0, 2, 1)                             # [N,V,4]
            plist.append(p[:, :, :3])
        p = torch.cat(plist, dim=1)                                                       # [N, sum V, 3]
        npts = p.shape[1]
        dist = mesh.distance(p.reshape(-1, 3).contiguous()).reshape(-1, npts)
        out.append(dist.min(dim=1).values)
    return torch.cat(out) if out else torch.zeros(0, device=dev)


def arm_append_list(X_list, Y_list, chain, link_vertices, triangles_obs, numsamples, dim, offset, margin,
                    link_boxes=None, obstacle_box=None):
    """Rejection sampler in joint space (:220-287): x in [-0.5, 0.5]^dim, configuration = (pi / 0.5) x, partner
    x1 = x0 + r u with r ~ U[0, sqrt(3)] (the reference keeps sqrt(3) for every dim, :231); keep pairs whose start
    configuration lies in the band offset <= distance <= margin (and passes the OBB filter when boxes are given)."""
    mesh = triangles_obs if isinstance(triangles_obs, ObstacleMesh) else ObstacleMesh(triangles_obs)
    dev = mesh.triangles.device
    OutsideSize = numsamples + 2
    WholeSize = 0
    scale = math.pi / 0.5
    while OutsideSize > 0:
        P = torch.rand((5 * numsamples, dim), dtype=torch.float32, device=dev) - 0.5
        dP = torch.rand((5 * numsamples, dim), dtype=torch.float32, device=dev) - 0.5
        rL = torch.rand((5 * numsamples, 1), dtype=torch.float32, device=dev) * np.sqrt(3)
        nP = P + torch.nn.functional.normalize(dP, dim=1) * rL
        inside = torch.all((nP <= 0.5), dim=1) & torch.all((nP >= -0.5), dim=1)
        x0, x1 = P[inside, :], nP[inside, :]
        if x0.shape[0] <= 1:
            continue
        if link_boxes is not None and obstacle_box is not None:
            where = arm_obstacle_obb(scale * x0, chain, link_boxes, obstacle_box, margin)
            x0, x1 = x0[where], x1[where]
            if x0.shape[0] == 0:
                continue
        d0 = arm_obstacle_distance(scale * x0, chain, link_vertices, mesh)
        where_d = (d0 >= offset) & (d0 <= margin)
        x0, x1, y0 = x0[where_d], x1[where_d], d0[where_d]
        y1 = arm_obstacle_distance(scale * x1, chain, link_vertices, mesh) if x1.shape[0] else y0
        X_list.append(torch.cat((x0, x1), 1))
        Y_list.append(torch.cat((y0.unsqueeze(1), y1.unsqueeze(1)), 1))
        OutsideSize = OutsideSize - x0.shape[0]
        WholeSize = WholeSize + x0.shape[0]
        if WholeSize > numsamples:
            break
    return X_list, Y_list


def arm_rand_sample_bound_points(numsamples, dim, v_obs, f_obs, offset, margin, chain, link_vertices,
                                 link_boxes=None, obstacle_box=None, device="cuda"):
    """numpy obstacle vertices / faces in, (sampled_points [n, 2*dim], speed [n, 2]) numpy out (:289-323)."""
    numsamples = int(numsamples)
    v_obs = torch.as_tensor(np.asarray(v_obs), dtype=torch.float32, device=device)
    f_obs = torch.as_tensor(np.asarray(f_obs), dtype=torch.long, device=device)
    mesh = ObstacleMesh(v_obs[f_obs].unsqueeze(dim=0))
    X_list, Y_list = arm_append_list([], [], chain, link_vertices, mesh, numsamples, dim, offset, margin, link_boxes,
                                     obstacle_box)
    X = torch.cat(X_list, 0)[:numsamples]
    Y = torch.cat(Y_list, 0)[:numsamples]
    sampled_points = X.detach().cpu().numpy()
    distance = Y.detach().cpu().numpy()
    speed = np.zeros((distance.shape[0], 2))
    speed[:, 0] = np.clip(distance[:, 0], a_min=offset, a_max=margin) / margin
    speed[:, 1] = np.clip(distance[:, 1], a_min=offset, a_max=margin) / margin
    return sampled_points, speed
