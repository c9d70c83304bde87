"""pntf_mesh_distance and the speed sampler (dataprocessing/speed_sampling_gpu.py:325-421) against
the oracle and closed forms.  Tolerance: fp32 distances in the unit box, absolute 2e-6."""
import numpy as np
import pytest
import torch

from oracle import mesh_oracle as MO

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-6


@pytest.fixture(scope="module")
def capi():
    from pntf import capi as c
    c.lib()
    return c


def _soup(rs, M, size=0.08):
    c = rs.uniform(-0.45, 0.45, size=(M, 1, 3))
    return (c + rs.uniform(-size, size, size=(M, 3, 3))).astype(np.float32)


@pytest.mark.parametrize("M,N", [(1, 1), (7, 300), (128, 257), (129, 1000), (1500, 5000)])
def test_mesh_distance_vs_oracle(capi, M, N):
    rs = np.random.RandomState(M + N)
    tri = _soup(rs, M)
    if M > 5:
        tri[3] = tri[3][0]                                  # a degenerate triangle (point)
        tri[4][2] = 0.5 * (tri[4][0] + tri[4][1])           # and a segment
    pts = rs.uniform(-0.5, 0.5, size=(N, 3)).astype(np.float32)
    want = MO.point_obstacle_distance(pts, tri)
    d, s = capi.mesh_distance(torch.tensor(tri, device=DEV), torch.tensor(pts, device=DEV), 0.004, 0.04)
    assert np.abs(d.cpu().numpy() - want).max() < TOL
    assert np.abs(s.cpu().numpy() - MO.speed_from_distance(want, 0.004, 0.04)).max() < 1e-4
    d_only = capi.mesh_distance(torch.tensor(tri, device=DEV), torch.tensor(pts, device=DEV))
    assert torch.equal(d_only, d)


def test_box_scene_full_size_closed_form_and_order_independence(capi):
    """1 Mi points against a scene of subdivided boxes: closed-form distances, and bit-identical results when
    triangles and points are shuffled (the kernel's tile skipping must never change the minimum)."""
    from pntf import speed_sampling as SS
    rs = np.random.RandomState(1)
    boxes = [(np.array([-0.3, -0.3, -0.3]), np.array([-0.05, 0.0, 0.1])),
             (np.array([0.1, -0.2, -0.4]), np.array([0.4, 0.35, -0.1])),
             (np.array([-0.1, 0.2, 0.15]), np.array([0.3, 0.45, 0.4]))]
    tri = np.concatenate([MO.box_mesh(lo, hi) for lo, hi in boxes])
    for _ in range(5):                                       # midpoint subdivision: 36 -> 36 864 triangles
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = (a + b) / 2, (b + c) / 2, (c + a) / 2
        tri = np.concatenate([np.stack(t, 1) for t in ((a, ab, ca), (ab, b, bc), (ca, bc, c), (ab, bc, ca))])
    N = 1 << 20
    pts = rs.uniform(-0.5, 0.5, size=(N, 3)).astype(np.float32)
    want = np.min([MO.box_distance(pts, lo, hi) for lo, hi in boxes], axis=0)
    t = torch.tensor(tri, dtype=torch.float32, device=DEV)
    p = torch.tensor(pts, device=DEV)
    d_raw = capi.mesh_distance(t, p)
    mesh = SS.ObstacleMesh(t.unsqueeze(0))
    d_sorted = mesh.distance(p)
    d_ref_api = SS.point_obstacle_distance(p, t.unsqueeze(0))
    perm_t = torch.randperm(t.shape[0], device=DEV)
    perm_p = torch.randperm(N, device=DEV)
    d_shuf = capi.mesh_distance(t[perm_t].contiguous(), p[perm_p].contiguous())
    torch.cuda.synchronize()
    assert np.abs(d_raw.cpu().numpy() - want).max() < TOL
    assert torch.equal(d_sorted, d_raw) and torch.equal(d_ref_api, d_raw)
    assert torch.equal(d_shuf, d_raw[perm_p])


def test_sampler_band_and_layout(capi):
    """point_rand_sample_bound_points: shapes, the start point lies in the band, both points inside the unit box,
    speed = clip(distance) / margin recomputed by the oracle, and a seeded run repeats."""
    from pntf import speed_sampling as SS
    v = np.array([[(0.2 if (k >> i) & 1 else -0.2) for i in range(3)] for k in range(8)])
    f = np.array([(0, 1, 3), (0, 3, 2), (4, 6, 7), (4, 7, 5), (0, 4, 5), (0, 5, 1), (2, 3, 7), (2, 7, 6), (0, 2, 6),
                  (0, 6, 4), (1, 5, 7), (1, 7, 3)])
    offset, margin = SS.band_for_task("test")
    torch.manual_seed(7)
    x, s = SS.point_rand_sample_bound_points(20000, 3, v, f, offset, margin, device=DEV)
    torch.manual_seed(7)
    x2, s2 = SS.point_rand_sample_bound_points(20000, 3, v, f, offset, margin, device=DEV)
    assert x.shape == (20000, 6) and s.shape == (20000, 2)
    assert np.array_equal(x, x2) and np.array_equal(s, s2)
    assert np.abs(x).max() <= 0.5
    d0 = MO.box_distance(x[:, :3], [-0.2] * 3, [0.2] * 3)
    d1 = MO.box_distance(x[:, 3:], [-0.2] * 3, [0.2] * 3)
    assert d0.min() > offset - TOL and d0.max() < margin + TOL
    assert np.abs(s[:, 0] - MO.speed_from_distance(d0, offset, margin)).max() < 1e-4
    assert np.abs(s[:, 1] - MO.speed_from_distance(d1, offset, margin)).max() < 1e-4


def test_sample_speed_writes_the_reference_files(capi, tmp_path):
    from pntf import speed_sampling as SS
    d = tmp_path / "datasets" / "test" / "scene"
    d.mkdir(parents=True)
    lines = ["OFF", "8 12 0"]
    lines += ["%g %g %g" % tuple((0.15 if (k >> i) & 1 else -0.25) for i in range(3)) for k in range(8)]
    for q in [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]:
        lines += ["3 %d %d %d" % (q[0], q[1], q[2]), "3 %d %d %d" % (q[0], q[2], q[3])]
    (d / "mesh_scaled.off").write_text("\n".join(lines) + "\n")
    SS.sample_speed(str(d / "mesh.obj"), 5000, 3, task_name="test", device=DEV)
    x, s, B = np.load(d / "sampled_points.npy"), np.load(d / "speed.npy"), np.load(d / "B.npy")
    assert x.shape == (5000, 6) and s.shape == (5000, 2) and B.shape == (3, 128)
    assert s.min() >= 0.1 - 1e-6 and s.max() <= 1.0 + 1e-6


def test_mesh_distance_rejects_bad_arguments(capi):
    t = torch.zeros(4, 3, 3, device=DEV)
    p = torch.zeros(5, 3, device=DEV)
    L = capi.lib()
    assert L.pntf_mesh_distance(t.data_ptr(), 4, p.data_ptr(), 5, None, 0.0, 0.0, None, None, 0, None) != 0
    out = torch.zeros(5, device=DEV)
    assert L.pntf_mesh_distance(t.data_ptr(), 4, p.data_ptr(), 5, out.data_ptr(), 0.0, 0.0, None, None, 0, None) != 0
    assert b"workspace" in L.pntf_last_error()
    assert L.pntf_mesh_distance(t.data_ptr(), 0, p.data_ptr(), 5, out.data_ptr(), 0.0, 0.0, None, None, 0, None) != 0


def _ur5_like_chain():
    """A six-joint serial chain with UR5-like offsets (synthetic: the UR5's URDF and meshes are not available)."""
    J = [("shoulder_link", (0, 0, 0.089), (0, 0, 0), (0, 0, 1)), ("upper_arm_link", (0, 0.136, 0), (0, 1.5708, 0), (0, 1, 0)),
         ("forearm_link", (0, -0.12, 0.425), (0, 0, 0), (0, 1, 0)), ("wrist_1_link", (0, 0, 0.392), (0, 1.5708, 0), (0, 1, 0)),
         ("wrist_2_link", (0, 0.093, 0), (0, 0, 0), (0, 0, 1)), ("wrist_3_link", (0, 0, 0.095), (0, 0, 0), (0, 1, 0))]
    return [{"name": n, "type": "revolute", "xyz": xyz, "rpy": rpy, "axis": ax} for n, xyz, rpy, ax in J]


def _urdf_of(joints):
    links = ["base_link"] + [j["name"] for j in joints]
    s = ["<robot name='synthetic'>"] + ["<link name='%s'/>" % l for l in links]
    for parent, j in zip(links, joints):
        s.append("<joint name='j_%s' type='%s'><parent link='%s'/><child link='%s'/><origin xyz='%s' rpy='%s'/>"
                 "<axis xyz='%s'/></joint>" % (j["name"], j["type"], parent, j["name"], " ".join(map(str, j["xyz"])),
                                               " ".join(map(str, j["rpy"])), " ".join(map(str, j["axis"]))))
    return "\n".join(s + ["</robot>"])


def test_arm_branch_fk_distance_and_obb_vs_oracle(capi):
    """Arm branch of the sampler (dataprocessing/speed_sampling_gpu.py:76-218): forward kinematics of a serial chain
    (built from URDF text), link vertices moved into the world, distance of the whole arm to the obstacle mesh with
    pntf_mesh_distance, and the OBB pre-filter -- against oracle/arm_oracle.py (numpy float64)."""
    from oracle import arm_oracle as AO
    from pntf import arm_sampling as AS
    joints = _ur5_like_chain()
    chain = AS.SerialChain.from_urdf(_urdf_of(joints), "wrist_3_link", device=DEV)
    assert chain.n_dof == 6 and chain.base_link == "base_link"
    rs = np.random.RandomState(4)
    th = rs.uniform(-np.pi, np.pi, size=(64, 6)).astype(np.float32)
    tg = chain.forward_kinematics(torch.from_numpy(th).to(DEV))
    assert list(tg.keys()) == ["base_link"] + [j["name"] for j in joints]
    for i in (0, 17, 63):
        for (name, T) in AO.forward_kinematics(joints, th[i])[1:]:
            np.testing.assert_allclose(tg[name][i].cpu().numpy(), T, atol=2e-6)
    boxes = {j["name"]: ((0.05, 0.05, 0.3), (-0.05, -0.05, 0.0)) for j in joints}
    verts = {n: AS.box_link_vertices(lo, hi, n=3) for n, (hi, lo) in boxes.items()}
    lo_o, hi_o = np.array([0.25, -0.3, 0.0]), np.array([0.6, 0.3, 0.5])
    tri = MO.box_mesh(lo_o, hi_o).astype(np.float32)
    d = AS.arm_obstacle_distance(torch.from_numpy(th).to(DEV), chain, verts, torch.from_numpy(tri).to(DEV).unsqueeze(0))
    want = AO.arm_obstacle_distance(th, joints, verts, tri, first_link=2)
    np.testing.assert_allclose(d.cpu().numpy(), want, atol=3e-6)
    assert want.min() < 0.02 < want.max()
    # OBB filter: the reference reports the LAST link's box only (:149); exact SAT oracle on that box
    link_boxes = [np.array(boxes[j["name"]][0] + boxes[j["name"]][1], np.float32) for j in joints[2:]]
    obs_box = np.concatenate([hi_o, lo_o]).astype(np.float32)
    margin = 0.04
    got = AS.arm_obstacle_obb(torch.from_numpy(th).to(DEV), chain, link_boxes, obs_box, margin).cpu().numpy()
    last = joints[-1]["name"]
    hi_b, lo_b = np.array(boxes[last][0]) + margin, np.array(boxes[last][1]) - margin
    for i in range(64):
        T = dict(AO.forward_kinematics(joints, th[i]))[last]
        assert got[i] == AO.obb_overlaps_aabb(T, hi_b, lo_b, hi_o, lo_o), i
    assert got.any() and not got.all()


def test_arm_sampler_band_and_layout(capi):
    from pntf import arm_sampling as AS
    joints = _ur5_like_chain()
    chain = AS.SerialChain(joints, device=DEV)
    verts = {j["name"]: AS.box_link_vertices((-0.05, -0.05, 0.0), (0.05, 0.05, 0.3)) for j in joints}
    lo_o, hi_o = np.array([0.25, -0.3, 0.0]), np.array([0.6, 0.3, 0.5])
    v, f = MO.box_mesh(lo_o, hi_o).reshape(-1, 3), np.arange(36).reshape(12, 3)
    torch.manual_seed(0)
    margin = 0.5 / 12
    pts, speed = AS.arm_rand_sample_bound_points(300, 6, v, f, margin / 10, margin, chain, verts, device=DEV)
    assert pts.shape == (300, 12) and speed.shape == (300, 2)
    assert np.abs(pts).max() <= 0.5 and speed.min() >= 0.1 - 1e-6 and speed.max() <= 1.0 + 1e-6
    d0 = AS.arm_obstacle_distance(torch.from_numpy(pts[:, :6]).to(DEV) * (np.pi / 0.5), chain, verts,
                                  torch.from_numpy(MO.box_mesh(lo_o, hi_o).astype(np.float32)).to(DEV).unsqueeze(0))
    assert float(d0.min()) >= margin / 10 - 1e-6 and float(d0.max()) <= margin + 1e-6      # starts lie in the band (:260)
    np.testing.assert_allclose(speed[:, 0], np.clip(d0.cpu().numpy(), margin / 10, margin) / margin, rtol=1e-5)
