"""GPU parity tests: the CUDA path (called through the C-ABI) against the committed golden
fixtures of the reference and against the CPU oracle on seeded inputs.  fp32 SIMT verify mode:
tau/speed/loss within rel 1e-5 (north_star), gradients within 5e-4 of autograd."""
import numpy as np
import pytest
import torch

from helpers import T, check_digest, load_golden, rel_err
from oracle import pntf_oracle as O

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def flat_from(P):
    return torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()


def cu(a):
    return T(np.asarray(a, np.float32)).to(DEV).contiguous()


@pytest.fixture(scope="module")
def capi():
    from pntf import capi as c
    c.lib()
    return c


def _pairs(rs, n, d):
    return rs.uniform(-0.5, 0.5, size=(n, 2 * d)).astype(np.float32)


@pytest.mark.parametrize("name,single", [("multi_d3.npz", False), ("single_d3.npz", True), ("single_d6.npz", True)])
def test_forward_modes_vs_golden(capi, name, single):
    fx = load_golden(name)
    d = int(fx["d"])
    P = O.init_params(int(fx["seed"]))
    flat = flat_from(P)
    if single:
        B = cu(fx["B"].T.copy()).view(1, d, 128)
        c = cu(fx["coords"])
    else:
        B = cu(fx["B"][:1])
        c = cu(fx["coords"][0])
    N = c.shape[0]
    tau0, _, _ = capi.forward(capi.MODE_VALUE, flat, B, c, 1, N, d)
    tau1, d1, _ = capi.forward(capi.MODE_GRAD, flat, B, c, 1, N, d)
    assert rel_err(tau0.cpu(), fx["out_tau"][:, 0]) < 1e-5
    assert rel_err(tau1.cpu(), fx["grad_tau"][:, 0]) < 1e-5
    assert rel_err(d1.cpu(), fx["grad_dtau"]) < 2e-5
    if single:
        tau2, d2, l2 = capi.forward(capi.MODE_LAPLACE, flat, B, c, 1, N, d)
        assert rel_err(tau2.cpu(), fx["lap_tau"][:, 0]) < 1e-5
        assert rel_err(d2.cpu(), fx["lap_dtau"]) < 2e-5
        assert rel_err(l2.cpu(), fx["lap_ltau"]) < 2e-4
    else:
        E = fx["coords"].shape[0]
        tau2, d2, l2 = capi.forward(capi.MODE_LAPLACE, flat, cu(fx["B"]), cu(fx["coords"]).view(E * N, 2 * d), E, N, d)
        assert rel_err(tau2.cpu().view(E, N), fx["lap_tau"][:, :, 0]) < 1e-5
        assert rel_err(d2.cpu().view(E, N, 2 * d), fx["lap_dtau"]) < 2e-5
        assert rel_err(l2.cpu().view(E, N, 2 * d), fx["lap_ltau"]) < 2e-4
        _, dc, _ = capi.forward(capi.MODE_GRAD, flat, B, c, 1, N, d, grad_flavour=capi.GRAD_COMPAT)
        assert rel_err(dc.cpu(), fx["backgrad_dtau"]) < 1e-4


@pytest.mark.parametrize("d,N", [(3, 1), (3, 37), (3, 1000), (6, 3), (6, 333)])
def test_forward_vs_oracle_ragged(capi, d, N):
    rs = np.random.RandomState(100 + d + N)
    P = O.init_params(21)
    Bw = (0.5 * rs.normal(size=(d, 128))).astype(np.float32)
    c = _pairs(rs, N, d)
    with torch.no_grad():
        tau, dtau, ltau = O.forward(P, T(c), T(Bw), 2)
    t2, d2, l2 = capi.forward(capi.MODE_LAPLACE, flat_from(P), cu(Bw).view(1, d, 128), cu(c), 1, N, d)
    assert rel_err(t2.cpu(), tau[:, 0]) < 1e-5
    assert rel_err(d2.cpu(), dtau) < 3e-5
    assert rel_err(l2.cpu(), ltau) < 3e-4


def test_forward_empty_batch(capi):
    P = O.init_params(21)
    tau, dtau, ltau = capi.forward(capi.MODE_LAPLACE, flat_from(P), torch.zeros(1, 3, 128, device=DEV),
                                   torch.zeros(0, 6, device=DEV), 1, 0, 3)
    assert tau.numel() == 0 and dtau.shape == (0, 6)


@pytest.mark.parametrize("name,variant", [("multi_d3.npz", "multi"), ("single_d3.npz", "single"), ("single_d6.npz", "single")])
def test_loss_and_param_grads_vs_golden(capi, name, variant):
    fx = load_golden(name)
    d = int(fx["d"])
    P = O.init_params(int(fx["seed"]))
    flat = flat_from(P)
    if variant == "single":
        B = cu(fx["B"].T.copy()).view(1, d, 128)
        E, N = 1, fx["coords"].shape[0]
        v = capi.LOSS_SINGLE
    else:
        B = cu(fx["B"])
        E, N = fx["coords"].shape[:2]
        v = capi.LOSS_MULTI
    c = cu(fx["coords"]).view(E * N, 2 * d)
    s = cu(fx["speed"]).view(E * N, 2)
    beta, gamma = float(fx["beta"]), float(fx["gamma"])
    lo, diff = capi.loss(v, flat, B, c, s, E, N, d, 1.0, beta, gamma)
    assert abs(lo[0].item() - fx["loss"]) / abs(fx["loss"]) < 1e-5
    assert abs(lo[1].item() - fx["loss_n"]) / abs(fx["loss_n"]) < 1e-5
    assert rel_err(diff.cpu().view(fx["diff"].shape), fx["diff"]) < 1e-4
    grad = torch.zeros_like(flat)
    lo2, diff2 = capi.loss(v, flat, B, c, s, E, N, d, 1.0, beta, gamma, grad_out=grad)
    assert abs(lo2[0].item() - fx["loss"]) / abs(fx["loss"]) < 1e-5
    assert torch.equal(diff2, diff) or rel_err(diff2.cpu(), diff.cpu()) < 1e-6
    from pntf.params import tensor_offsets, tensor_shapes
    offs, shapes = tensor_offsets(), tensor_shapes()
    g = grad.cpu().numpy()
    for k in offs:
        n = int(np.prod(shapes[k]))
        gk = g[offs[k]:offs[k] + n].reshape(shapes[k])
        if k.startswith("encoder1.0."):
            assert np.all(gk == 0)
            continue
        check_digest(fx, "grad/", k, gk, 1e-3)


@pytest.mark.parametrize("d,variant,N", [(3, "multi", 203), (6, "single", 50)])
def test_param_grads_vs_oracle_autograd_chunked(capi, d, variant, N):
    """Ragged N, tiny stash budget (forces several chunks inside one pntf_loss_bwd call), alpha != 1."""
    rs = np.random.RandomState(7 + d)
    P = O.init_params(5)
    E = 2 if variant == "multi" else 1
    Bw = (0.5 * rs.normal(size=(E, d, 128))).astype(np.float32)
    c = np.stack([_pairs(rs, N, d) for _ in range(E)])
    s = rs.uniform(0.1, 1.0, size=(E, N, 2)).astype(np.float32)
    alpha, beta, gamma = 0.8, 2.5, 1e-3
    s_int = alpha * s + 1 - alpha
    if variant == "multi":
        loss, loss_n, diff, grads = O.loss_and_grads(P, T(c), T(s_int), T(Bw), beta, gamma, "multi")
    else:
        loss, loss_n, diff, grads = O.loss_and_grads(P, T(c[0]), T(s_int[0]), T(Bw[0]), beta, gamma, "single")
    flat = flat_from(P)
    grad = torch.zeros_like(flat)
    v = capi.LOSS_MULTI if variant == "multi" else capi.LOSS_SINGLE
    per_pair = capi.workspace_bytes(capi.OP_LOSS_BWD, 1, 64, d) // 64
    budget = int(per_pair * 48)        # ~48 pairs per chunk
    lo, df = capi.loss(v, flat, cu(Bw), cu(c).view(E * N, 2 * d), cu(s).view(E * N, 2), E, N, d, alpha, beta, gamma,
                       grad_out=grad, budget=budget)
    assert abs(lo[0].item() - loss.item()) / abs(loss.item()) < 1e-5
    assert rel_err(df.cpu(), diff.reshape(-1)) < 1e-4
    from pntf.params import tensor_offsets, tensor_shapes
    offs, shapes = tensor_offsets(), tensor_shapes()
    g = grad.cpu()
    for k, ref in grads.items():
        n = int(np.prod(shapes[k]))
        gk = g[offs[k]:offs[k] + n].view(shapes[k])
        if ref is None:
            assert bool((gk == 0).all())
        else:
            assert rel_err(gk, ref) < 1e-3, (k, rel_err(gk, ref))


def test_fields_and_plan_vs_golden(capi):
    fx = load_golden("multi_d3.npz")
    d = 3
    P = O.init_params(int(fx["seed"]))
    flat = flat_from(P)
    B = cu(fx["B"][:1])
    c = cu(fx["coords"][0])
    N = c.shape[0]
    tt = capi.field(capi.FIELD_TRAVEL_TIME, flat, B, c, 1, N, d)
    sp = capi.field(capi.FIELD_SPEED, flat, B, c, 1, N, d)
    gr = capi.field(capi.FIELD_DIRECTION, flat, B, c, 1, N, d, grad_flavour=capi.GRAD_COMPAT)
    assert rel_err(tt.cpu(), fx["travel_times"]) < 1e-5
    assert rel_err(sp.cpu(), fx["speed_pred"]) < 1e-5 * 10
    assert rel_err(gr.cpu(), fx["gradient"]) < 1e-4
    way, nst = capi.plan(flat, B, cu(fx["plan_start_goal"]), d, float(fx["plan_step"]), float(fx["plan_thr"]),
                         int(fx["plan_cap"]), capi.GRAD_COMPAT)
    assert np.array_equal(nst.cpu().numpy(), fx["plan_nsteps"])
    L = fx["plan_waypoints"].shape[1]
    assert np.abs(way.cpu().numpy()[:, :L] - fx["plan_waypoints"]).max() < 1e-3
    # rows after the last step repeat the final position
    w = way.cpu().numpy()
    for q in range(w.shape[0]):
        assert np.all(w[q, nst[q].item():] == w[q, nst[q].item()])


@pytest.mark.parametrize("name", ["single_d3.npz", "single_d6.npz"])
def test_single_fields_and_plan_vs_golden(capi, name):
    fx = load_golden(name)
    d = int(fx["d"])
    P = O.init_params(int(fx["seed"]))
    flat = flat_from(P)
    B = cu(fx["B"].T.copy()).view(1, d, 128)
    c = cu(fx["coords"])
    N = c.shape[0]
    assert rel_err(capi.field(capi.FIELD_TRAVEL_TIME, flat, B, c, 1, N, d).cpu(), fx["travel_times"]) < 1e-5
    assert rel_err(capi.field(capi.FIELD_SPEED, flat, B, c, 1, N, d).cpu(), fx["speed_pred"]) < 1e-4
    assert rel_err(capi.field(capi.FIELD_SPEED2, flat, B, c, 1, N, d, gamma=float(fx["gamma"])).cpu(), fx["speed2_pred"]) < 1e-4
    n = fx["gradient_rows"].shape[0]
    assert rel_err(capi.field(capi.FIELD_DIRECTION, flat, B, c[:n].contiguous(), 1, n, d).cpu(), fx["gradient_rows"]) < 1e-4
    way, nst = capi.plan(flat, B, cu(fx["plan_start_goal"]), d, float(fx["plan_step"]), float(fx["plan_thr"]),
                         int(fx["plan_cap"]), capi.GRAD_EXACT)
    assert np.array_equal(nst.cpu().numpy(), fx["plan_nsteps"])
    L = fx["plan_waypoints"].shape[1]
    assert np.abs(way.cpu().numpy()[:, :L] - fx["plan_waypoints"]).max() < 1e-3


def test_plan_batched_vs_oracle_and_cap(capi):
    d = 3
    rs = np.random.RandomState(9)
    P = O.init_params(8)
    Bw = (0.5 * rs.normal(size=(d, 128))).astype(np.float32)
    sg = rs.uniform(-0.4, 0.4, size=(19, 2 * d)).astype(np.float32)
    sg[3, d:] = sg[3, :d] + 0.01          # already within thr: zero steps
    with torch.no_grad():
        way, nst = O.plan(P, T(sg), T(Bw), 0.03, 0.06, 12, "exact")     # cap hit by the long ones
    w2, n2 = capi.plan(flat_from(P), cu(Bw).view(1, d, 128), cu(sg), d, 0.03, 0.06, 12, capi.GRAD_EXACT)
    assert np.array_equal(n2.cpu().numpy(), nst.numpy())
    assert n2[3].item() == 0 and int(n2.max()) == 13
    assert np.abs(w2.cpu().numpy() - way.numpy()).max() < 1e-3


def test_adamw_known_answer(capi):
    fx = load_golden("adamw_kat.npz")
    p = cu(fx["p0"]).clone()
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)
    mask = torch.ones(p.numel(), dtype=torch.uint8, device=DEV)
    mask[:16] = 0
    for t in range(3):
        capi.adamw(p, cu(fx["grads"][t]), m, v, mask, t + 1, lr=float(fx["lrs"][t]))
        ref = fx["p%d" % (t + 1)]
        assert np.abs(p.cpu().numpy()[16:] - ref[16:]).max() < 2e-7
        assert np.array_equal(p.cpu().numpy()[:16], fx["p0"][:16])      # masked: untouched, no decay
    assert rel_err(m.cpu()[16:], fx["m3"][16:]) < 1e-6 and rel_err(v.cpu()[16:], fx["v3"][16:]) < 1e-6


def test_reference_shaped_api_and_autograd(capi):
    from models import model_res_sigmoid_multi as M, model_res_sigmoid as S
    fx = load_golden("multi_d3.npz")
    d = 3
    model = M.Model("/tmp", "/tmp", d, 2, device=DEV)
    net = M.NN(DEV, d)
    net.load_state_dict(O.init_params(int(fx["seed"])), strict=True)
    net.to(DEV)
    net.precision = "fp32"          # verify mode: the golden tolerances below are the fp32 ones
    model.network = net
    c, s, B = cu(fx["coords"]), cu(fx["speed"]), cu(fx["B"])
    tau, dtau, ltau, Xp = net.out_laplace(c, B)
    assert tau.shape == (2, 40, 1) and dtau.shape == (2, 40, 6) and Xp is c
    assert rel_err(ltau.cpu(), fx["lap_ltau"]) < 2e-4
    tb, db, _ = net.out_backgrad(c[0], B[0])
    assert rel_err(db.cpu(), fx["backgrad_dtau"]) < 1e-4
    loss, loss_n, diff = model.Loss(c, s, B, float(fx["beta"]), float(fx["gamma"]))
    assert diff.shape == (2, 40)
    loss.backward()
    named = dict(net.named_parameters())
    assert named["encoder1.0.weight"].grad is None and named["encoder1.0.bias"].grad is None
    for k, p in named.items():
        if p.grad is not None:
            check_digest(fx, "grad/", k, p.grad.cpu().numpy(), 1e-3)
    # torch's own AdamW on the (view) parameters reproduces the reference's first step
    opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=0.1)
    opt.step()
    for k, p in net.state_dict().items():
        check_digest(fx, "adamw/", k, p.cpu().numpy(), 1e-4)
    model.B = B[0]
    net.load_state_dict(O.init_params(int(fx["seed"])), strict=True)
    assert rel_err(model.TravelTimes(c[0]).cpu(), fx["travel_times"]) < 1e-5
    assert rel_err(model.Gradient(c[0], B[0]).cpu(), fx["gradient"]) < 1e-4
    way, nst = model.plan(cu(fx["plan_start_goal"]), B[0], 0.03, 0.06, 500)
    assert np.array_equal(nst.cpu().numpy(), fx["plan_nsteps"])
    path = model.path(way, nst, 0, d)
    assert path.shape == (2 * (int(nst[0]) + 1), d)

    fs = load_golden("single_d6.npz")
    ms = S.Model("/tmp", "/tmp", 6, device=DEV)
    ns = S.NN(DEV, 6, T(fs["B"]))
    ns.load_state_dict(O.init_params(int(fs["seed"])), strict=True)
    ns.to(DEV)
    ns.precision = "fp32"
    ms.network = ns
    loss, loss_n, diff = ms.Loss(cu(fs["coords"]), cu(fs["speed"]), float(fs["beta"]), float(fs["gamma"]))
    assert abs(loss.item() - fs["loss"]) / abs(fs["loss"]) < 1e-5
    assert rel_err(ms.Speed2(cu(fs["coords"]), float(fs["gamma"])).cpu(), fs["speed2_pred"]) < 1e-4


def test_training_loop_reduces_loss(capi):
    """Short run of the trainer control flow (alpha ramp, beta rule, fused AdamW) on synthetic data."""
    from models import model_res_sigmoid_multi as M
    from pntf import synth
    torch.manual_seed(0)
    d = 3
    envs = []
    for e in range(2):
        c, s = synth.sample_pairs(10000, d, DEV, seed=10 + e, margin=0.5 / 5)     # one step of 2 x 10 000 pairs per epoch
        envs.append((torch.cat([c, s], dim=1), synth.sample_B(d, DEV, seed=20 + e)[0]))
    model = M.Model("", "", d, 2, device=DEV)
    hist = model.train(envs=envs, epochs=6, log=lambda *a: None)
    assert len(hist) == 6
    assert all(np.isfinite(h[1]) for h in hist)
    assert hist[-1][1] < hist[0][1]


def test_tc_selftest_plumbing(capi):
    """tcgen05 operand layouts / descriptors / TMEM mapping against numpy (fp16-rounded operands)."""
    rs = np.random.RandomState(5)
    W = rs.normal(size=(128, 128)).astype(np.float32)
    X = rs.normal(size=(128, 128)).astype(np.float32)
    C = rs.normal(size=(128, 128)).astype(np.float32)
    out = capi.tc_selftest(cu(W), cu(X), cu(C)).cpu().numpy()
    Wh = W.astype(np.float16).astype(np.float64)
    Xh = X.astype(np.float16).astype(np.float64)
    assert np.abs(out[0] - Wh @ Xh).max() < 2e-4
    assert np.abs(out[1] - (C + Wh @ Xh)).max() < 2e-4
    assert np.abs(out[2] - Xh @ Xh.T).max() < 2e-4
