"""TEST INFRASTRUCTURE ONLY (never imported by the product path): CPU restatement of the arm branch of the ground-truth
speed sampling, dataprocessing/speed_sampling_gpu.py:153-218 (arm_obstacle_distance) and :76-151 (arm_obstacle_obb).

  per configuration th:  T_0 = I;  T_k = T_{k-1} . Origin_k . Motion_k(th_k)   (URDF joint: origin xyz / rpy, then a rotation
  about `axis` for revolute joints or a translation along it for prismatic ones; what pytorch_kinematics' serial chain
  evaluates at :160-163);  every vertex v of link k's collision mesh goes to T_k [v 1]^T (:173-178);  distance of all of
  them to the obstacle triangles (:197, third-party bvh_distance_queries: here oracle/mesh_oracle.py, brute force in
  float64);  minimum over the arm (:203-204).

Parity status: UNPINNED against the reference binary -- pytorch_kinematics, bvh_distance_queries, igl and the UR5 files are
absent, the reference function cannot run here.  Pinned to closed forms instead (tests/test_cpu_mesh_oracle.py: a planar
two-link arm whose end point position is cos / sin of the joint angles, distance to a wall).
"""
import math

import numpy as np

from . import mesh_oracle as MO


def rpy_matrix(rpy):
    r, p, y = rpy
    Rx = np.array([[1, 0, 0], [0, math.cos(r), -math.sin(r)], [0, math.sin(r), math.cos(r)]])
    Ry = np.array([[math.cos(p), 0, math.sin(p)], [0, 1, 0], [-math.sin(p), 0, math.cos(p)]])
    Rz = np.array([[math.cos(y), -math.sin(y), 0], [math.sin(y), math.cos(y), 0], [0, 0, 1]])
    return Rz @ Ry @ Rx                      # URDF fixed-axis roll, pitch, yaw


def axis_angle(axis, q):
    a = np.asarray(axis, np.float64)
    a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + math.sin(q) * K + (1 - math.cos(q)) * (K @ K)


def forward_kinematics(joints, th):
    """joints: list of {name, type, xyz, rpy, axis}; th [n_dof] -> [(link name, T[4,4])] incl. the base (identity)."""
    T = np.eye(4)
    out = [("base", T.copy())]
    k = 0
    for j in joints:
        O = np.eye(4)
        O[:3, :3] = rpy_matrix(j.get("rpy", (0, 0, 0)))
        O[:3, 3] = j.get("xyz", (0, 0, 0))
        T = T @ O
        if j["type"] != "fixed":
            M = np.eye(4)
            if j["type"] == "prismatic":
                a = np.asarray(j["axis"], np.float64)
                M[:3, 3] = th[k] * a / np.linalg.norm(a)
            else:
                M[:3, :3] = axis_angle(j["axis"], th[k])
            k += 1
            T = T @ M
        out.append((j["name"], T.copy()))
    return out


def arm_obstacle_distance(th_batch, joints, link_vertices, triangles, first_link=2):
    res = []
    for th in np.asarray(th_batch, np.float64):
        pts = []
        for it, (name, T) in enumerate(forward_kinematics(joints, th)):
            if it < first_link or name not in link_vertices:
                continue
            v = np.asarray(link_vertices[name], np.float64)
            pts.append(v @ T[:3, :3].T + T[:3, 3])
        res.append(MO.point_obstacle_distance(np.concatenate(pts), triangles).min())
    return np.asarray(res)


def obb_overlaps_aabb(T, box_max, box_min, obs_max, obs_min):
    """Exact separating-axis test (15 axes) of the box [box_min, box_max] moved by T against the world box."""
    c = np.array([[(box_max if (k >> i) & 1 else box_min)[i] for i in range(3)] for k in range(8)])
    rob = c @ T[:3, :3].T + T[:3, 3]
    obs = np.array([[(obs_max if (k >> i) & 1 else obs_min)[i] for i in range(3)] for k in range(8)])
    R = T[:3, :3]
    axes = list(np.eye(3)) + [R[i, :] for i in range(3)] + [np.cross(R[i, :], e) for i in range(3) for e in np.eye(3)]
    for ax in axes:
        ra, oa = rob @ ax, obs @ ax
        if ra.max() < oa.min() or oa.max() < ra.min():
            return False
    return True
