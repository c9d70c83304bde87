"""The mesh-distance oracle (oracle/mesh_oracle.py) against closed forms, and the host-side pieces of the speed
sampler that need no GPU (dataprocessing/speed_sampling_gpu.py:325-336, 415-419, 457-472)."""
import numpy as np

from oracle import mesh_oracle as MO


def test_single_triangle_regions():
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], float)
    pts = np.array([[0.25, 0.25, 0.5],      # face
                    [-1, -1, 0],            # vertex a
                    [2, -0.5, 0],           # vertex b
                    [-0.5, 2, 0],           # vertex c
                    [0.5, -1, 0],           # edge ab
                    [-1, 0.5, 0],           # edge ac
                    [1, 1, 0],              # edge bc
                    [0.25, 0.25, 0.0]])     # on the triangle
    want = np.array([0.5, np.sqrt(2), np.sqrt(1.25), np.sqrt(1.25), 1.0, 1.0, np.sqrt(0.5), 0.0])
    got = MO.point_obstacle_distance(pts, tri)
    assert np.abs(got - want).max() < 1e-12


def test_box_mesh_matches_closed_form():
    rs = np.random.RandomState(0)
    lo, hi = np.array([-0.2, -0.1, -0.3]), np.array([0.25, 0.3, 0.1])
    pts = rs.uniform(-0.5, 0.5, size=(4000, 3))
    got = MO.point_obstacle_distance(pts, MO.box_mesh(lo, hi))
    assert np.abs(got - MO.box_distance(pts, lo, hi)).max() < 1e-12


def test_degenerate_triangles_are_harmless():
    tri = np.array([[[0, 0, 0], [0, 0, 0], [0, 0, 0]],            # a point
                    [[0, 0, 1], [1, 0, 1], [2, 0, 1]]], float)    # a segment
    pts = np.array([[0, 0, 0.25], [1.5, 0, 2.0]])
    got = MO.point_obstacle_distance(pts, tri)
    assert np.allclose(got, [0.25, 1.0])


def test_speed_clip():
    d = np.array([0.0, 0.001, 0.02, 0.05])
    s = MO.speed_from_distance(d, 0.004, 0.04)
    assert np.allclose(s, [0.1, 0.1, 0.5, 1.0])


def test_off_reader_and_band(tmp_path):
    from pntf import speed_sampling as SS
    p = tmp_path / "m_scaled.off"
    p.write_text("OFF\n4 2 0\n0 0 0\n1 0 0\n1 1 0\n0 1 0\n3 0 1 2\n4 0 1 2 3\n")
    v, f = SS.read_off(str(p))
    assert v.shape == (4, 3) and f.tolist() == [[0, 1, 2], [0, 1, 2], [0, 2, 3]]
    assert np.allclose(SS.band_for_task("gibson"), (0.5 / 120, 0.5 / 12))
    assert np.allclose(SS.band_for_task("test"), (0.01, 0.1))


def test_arm_oracle_planar_two_link_closed_form():
    """oracle/arm_oracle.py against a closed form: a planar two-link arm (links of length 1 along x, joints about z) and
    a wall x = 3: the end point is (cos a + cos(a+b), sin a + sin(a+b)), the arm's distance to the wall is 3 - max x."""
    from oracle import arm_oracle as AO
    joints = [{"name": "l1", "type": "revolute", "xyz": (0, 0, 0), "rpy": (0, 0, 0), "axis": (0, 0, 1)},
              {"name": "l2", "type": "revolute", "xyz": (1, 0, 0), "rpy": (0, 0, 0), "axis": (0, 0, 1)},
              {"name": "tip", "type": "fixed", "xyz": (1, 0, 0), "rpy": (0, 0, 0), "axis": (1, 0, 0)}]
    seg = np.stack([np.linspace(0, 1, 21), np.zeros(21), np.zeros(21)], axis=1)
    verts = {"l1": seg, "l2": seg, "tip": np.zeros((1, 3))}
    wall = np.array([[[3, -50, -50], [3, 50, -50], [3, 0, 80]]], np.float64)       # one huge triangle in the plane x = 3
    rs = np.random.RandomState(0)
    th = rs.uniform(-np.pi, np.pi, size=(40, 2))
    fk = [AO.forward_kinematics(joints, t) for t in th]
    tips = np.array([f[-1][1][:3, 3] for f in fk])
    np.testing.assert_allclose(tips[:, 0], np.cos(th[:, 0]) + np.cos(th[:, 0] + th[:, 1]), atol=1e-12)
    np.testing.assert_allclose(tips[:, 1], np.sin(th[:, 0]) + np.sin(th[:, 0] + th[:, 1]), atol=1e-12)
    d = AO.arm_obstacle_distance(th, joints, verts, wall, first_link=1)
    elbow_x = np.cos(th[:, 0])
    want = 3.0 - np.maximum.reduce([np.zeros(40), elbow_x, tips[:, 0]])
    np.testing.assert_allclose(d, want, atol=1e-9)
    # rpy convention: yaw about z after pitch after roll (URDF fixed axes)
    R = AO.rpy_matrix((0.3, -0.2, 0.7))
    np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-12)
    np.testing.assert_allclose(AO.rpy_matrix((0, 0, np.pi / 2)) @ [1, 0, 0], [0, 1, 0], atol=1e-12)
    # separating-axis oracle: a unit cube rotated 45 degrees about z touches / misses a neighbour as geometry says
    T = np.eye(4); T[:3, :3] = AO.axis_angle((0, 0, 1), np.pi / 4); T[:3, 3] = (2.0, 0, 0)
    assert AO.obb_overlaps_aabb(T, (0.5,) * 3, (-0.5,) * 3, (1.4, 1, 1), (0, -1, -1))          # corner reaches x = 2 - 0.707
    assert not AO.obb_overlaps_aabb(T, (0.5,) * 3, (-0.5,) * 3, (1.2, 1, 1), (0, -1, -1))
