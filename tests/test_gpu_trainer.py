"""GPU tests of the host-side rows through the reference-shaped API: step-by-step replay of recorded runs of the
UNMODIFIED reference trainers (T2; tests/golden/make_golden_train.py), the rollback branch, checkpoint load -> inference,
field slices against the arrays Model.plot draws (8f rank 4), and the autograd / optimiser semantics of the drop-in."""
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN, load_golden, train_fixture_env, write_multi_dataset, write_single_dataset

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def capi():
    from pntf import capi as c
    c.lib()
    return c


def _load_params(net, P):
    fp = net.flat()
    for k, v in P.items():
        fp.views[k].copy_(v.to(fp.device))


class ReplayHooks:
    """Feeds the permutations the reference drew (in its order) and records the schedule."""

    def __init__(self, env_orders, row_perms, rows, randints=()):
        from pntf.trainer import TrainHooks  # noqa: F401  (interface documented there)
        self.env_orders = [torch.from_numpy(np.asarray(e, np.int64)) for e in env_orders]
        self.row_perms = list(row_perms)
        self.rows = rows
        self.randints = list(randints)
        self.steps, self.epochs = [], []

    def randperm(self, n):
        if n != self.rows:
            return self.env_orders.pop(0)
        head = np.asarray(self.row_perms.pop(0), np.int64)
        if head.size < n:     # the fixture keeps the rows a step can use; the rest of the permutation is never read
            rest = np.setdiff1d(np.arange(n, dtype=np.int64), head, assume_unique=False)
            head = np.concatenate([head, rest])
        return torch.from_numpy(head)

    def rollback_index(self):
        return self.randints.pop(0)

    def on_step(self, rec):
        self.steps.append({k: (v.detach().cpu().numpy().copy() if torch.is_tensor(v) and k == "loss_out" else v)
                           for k, v in rec.items() if k != "index"})

    def on_epoch(self, rec):
        self.epochs.append(dict(rec))


@pytest.mark.parametrize("prec,tol_first,tol_run", [("fp32", 2e-5, 1e-2), ("tc", 1e-2, 3e-2)])
def test_multi_trainer_replays_the_recorded_reference_run(capi, tmp_path, prec, tol_first, tol_run):
    """models/model_res_sigmoid_multi.py:953-1141 run unmodified on CPU for 8 epochs (2 environments x 12 000 rows,
    one 2 x 10 000-pair step per epoch); the same weights, permutations and data through Model.train's loop here.
    Step 1 must agree to kernel accuracy; later steps drift because AdamW's first updates are +-lr * sign(g) (tiny
    gradients may flip): measured 1e-6 for the first five epochs and 5e-3 by epoch 8 in fp32 mode."""
    from oracle import pntf_oracle as O
    from pntf import data as pdata
    from pntf.model import Model_multi
    from pntf.trainer import run_training
    fx = load_golden("train_multi_d3.npz")
    seed, d, rows, epochs = int(fx["seed"]), int(fx["d"]), int(fx["rows"]), int(fx["epochs"])
    root = str(tmp_path / "data") + "/"
    write_multi_dataset(root, [train_fixture_env(seed + i, rows, d) for i in range(2)])
    envs = pdata.load_multi(root, 2)
    model = Model_multi(str(tmp_path), root, d, 2, device=DEV)
    model.init_network()
    model.network.precision = prec
    _load_params(model.network, O.init_params(seed))
    hooks = ReplayHooks(fx["env_orders"], fx["row_perms"], rows)
    logs = []
    run_training(model, [e[0] for e in envs], [e[1] for e in envs], epochs=epochs, log=logs.append, hooks=hooks)
    ref = fx["steps"]                    # beta, gamma, lr, loss, loss_n, first speed value
    assert len(hooks.steps) == len(ref) == epochs
    got_loss_n = np.array([s["loss_out"][1] for s in hooks.steps], np.float64)
    got_loss = np.array([s["loss_out"][0] for s in hooks.steps], np.float64)
    assert abs(got_loss_n[0] - ref[0, 4]) / ref[0, 4] < tol_first
    if prec == "fp32":       # measured: 6 matching digits for the first five epochs, 5e-3 by epoch 8 (chaotic drift)
        np.testing.assert_allclose(got_loss_n[:5], ref[:5, 4], rtol=1e-4)
    np.testing.assert_allclose(got_loss_n, ref[:, 4], rtol=tol_run)
    np.testing.assert_allclose(got_loss, ref[:, 3], rtol=tol_run)
    np.testing.assert_allclose([s["beta"] for s in hooks.steps], ref[:, 0], rtol=tol_run)
    assert [s["gamma"] for s in hooks.steps] == list(ref[:, 1])
    assert [s["lr"] for s in hooks.steps] == list(ref[:, 2])
    np.testing.assert_allclose([e["alpha"] for e in hooks.epochs], fx["epoch_log"][:, 2])
    np.testing.assert_allclose([e["total_diff"] for e in hooks.epochs], fx["epoch_log"][:, 1], rtol=max(tol_run, 1e-4))
    np.testing.assert_allclose(model.total_train_loss, fx["total_train_loss"], rtol=tol_run)
    # same files as the reference wrote, modulo the .jpg plots (headless: .npz) and the loss digits in the names
    files = sorted(os.listdir(str(tmp_path)))
    for ep in (1, epochs):
        assert sum(f.startswith("Model_Epoch_%05d_ValLoss_" % ep) and f.endswith(".pt") for f in files) == 1
    assert sum(f.endswith(".pt") for f in files) == sum(str(f).endswith(".pt") for f in fx["files"])
    # M:1069: self.B is the first environment of the last batch
    np.testing.assert_array_equal(model.B.cpu().numpy(), fx["final_B"])
    # final weights: Frobenius norms per tensor
    sd = model.network.state_dict()
    for k, v in sd.items():
        fro = float(torch.linalg.norm(v.double()))
        assert abs(fro - fx["final/" + k + "#fro"]) / fx["final/" + k + "#fro"] < (2e-3 if prec == "fp32" else 1e-2), k
    assert all("Epoch = " in line for line in logs) and len(logs) == epochs


def test_single_trainer_replays_the_recorded_reference_run(capi, tmp_path):
    """models/model_res_sigmoid.py:938-1137 (d = 6, 12 000 rows: steps of 10 000 and 2 000 pairs per epoch; the loader
    shuffles in place, so epoch e sees the composition of the first e permutations)."""
    from oracle import pntf_oracle as O
    from pntf import data as pdata
    from pntf.model import Model_single
    from pntf.trainer import run_training
    fx = load_golden("train_single_d6.npz")
    seed, d, rows, epochs = int(fx["seed"]), int(fx["d"]), int(fx["rows"]), int(fx["epochs"])
    pts, spd, _ = train_fixture_env(seed, rows, d)
    write_single_dataset(str(tmp_path / "data"), pts, spd)
    data = pdata.load_single(str(tmp_path / "data"))
    model = Model_single(str(tmp_path), str(tmp_path / "data"), d, device=DEV)
    model.init_network(B=torch.from_numpy(fx["B"]))
    model.network.precision = "fp32"
    _load_params(model.network, O.init_params(seed))
    hooks = ReplayHooks([], fx["row_perms"], rows)
    run_training(model, [data], [model.network.B], epochs=epochs, log=lambda s: None, hooks=hooks)
    ref = fx["steps"]
    assert len(hooks.steps) == len(ref) == 2 * epochs
    got = np.array([s["loss_out"][1] for s in hooks.steps], np.float64)
    assert abs(got[0] - ref[0, 4]) / ref[0, 4] < 2e-5
    np.testing.assert_allclose(got, ref[:, 4], rtol=5e-3)
    np.testing.assert_allclose([s["beta"] for s in hooks.steps], ref[:, 0], rtol=5e-3)
    # S:1083-1084: / len(dataloader) = ceil(12000 / 10000) = 2
    np.testing.assert_allclose([e["total_diff"] for e in hooks.epochs], fx["epoch_log"][:, 1], rtol=5e-3)
    np.testing.assert_allclose(model.total_train_loss, fx["total_train_loss"], rtol=5e-3)


def test_rollback_restores_the_chosen_snapshot(capi, tmp_path):
    """S:1089-1097 / M:1100-1109: a loss ratio outside (0, 1.2) reloads network AND optimiser from one of the snapshots
    taken at the start of the previous <= 5 epochs and redoes the epoch; lr and beta are not reset."""
    from pntf.model import Model_multi
    from pntf.trainer import run_training
    rows, d = 10000, 3
    envs = [train_fixture_env(70 + i, rows, d) for i in range(2)]
    datas = [torch.from_numpy(np.concatenate([p.astype(np.float16).astype(np.float32), s], 1)) for p, s, _ in envs]
    Bs = [torch.from_numpy(b) for _, _, b in envs]
    torch.manual_seed(1)
    model = Model_multi("", "", d, 2, device=DEV)
    model.init_network()
    fp, opt = model.network.flat(), model.optimizer
    states, betas = [], []
    orig = model._loss_step_gather
    spikes = {0: 50.0, 4: 50.0}       # call 0 = epoch 1, first attempt; call 4 = epoch 4, first attempt (call 1 redoes epoch 1)

    def wrapped(table, index, B, E, N, alpha, beta, gamma):
        c = len(states)
        states.append((fp.flat.clone(), opt.exp_avg.clone(), opt.exp_avg_sq.clone(), opt.step_count))
        betas.append(beta)
        out = orig(table, index, B, E, N, alpha, beta, gamma)
        if c in spikes:
            fp.loss_slot.mul_(spikes[c])
        return out

    model._loss_step_gather = wrapped

    class Hooks:
        def __init__(self):
            self.gen = torch.Generator().manual_seed(0)
            self.picks = [4, 3]        # epoch 1: only one snapshot exists -> clamped to it; epoch 4: ring = [S0, S0, S1, S2]
            self.epochs = []

        def randperm(self, n):
            return torch.randperm(n, generator=self.gen)

        def rollback_index(self):
            return self.picks.pop(0)

        def on_step(self, rec):
            pass

        def on_epoch(self, rec):
            self.epochs.append(rec)

    hooks = Hooks()
    logs = []
    run_training(model, datas, Bs, epochs=5, log=logs.append, hooks=hooks, graph=False)   # launch by launch: the wrapper sees every step

    def same(a, b):
        return all(torch.equal(x, y) for x, y in zip(a[:3], b[:3])) and a[3] == b[3]

    assert len(states) == 7 and not hooks.picks
    assert same(states[1], states[0])               # epoch 1 redone from the initial state
    assert not same(states[2], states[1])
    # calls: 0,1 epoch 1 | 2 epoch 2 | 3 epoch 3 | 4,5 epoch 4 | 6 epoch 5 ; pick 3 of [S0, S0, S1, S2] = start of epoch 3
    assert same(states[5], states[3]) and states[5][3] == 2
    assert not same(states[5], states[4])
    assert betas[5] == betas[4] and betas[1] == betas[0] == 1.0
    assert [e["repeats"] for e in hooks.epochs] == [1, 0, 0, 1, 0]
    assert sum("RepeatEpoch" in line for line in logs) == 2


def test_reference_checkpoint_to_inference_and_field_slices(capi, tmp_path):
    """Loads the reference-written checkpoint fixture and evaluates through the public API; field_slice against the
    arrays the reference's Model.plot hands to matplotlib (M:1250-1293, S:1284-1329)."""
    from oracle import pntf_oracle as O
    from pntf.model import Model_multi, Model_single
    ck = os.path.join(GOLDEN, "ref_ckpt", "Model_Epoch_00003_ValLoss_1.250000e-02.pt")
    m = Model_multi(str(tmp_path), "", 3, 2, device=DEV)
    m.load(ck)
    m.B = torch.load(ck, map_location="cpu", weights_only=False)["B_state_dict"].to(DEV)
    rs = np.random.RandomState(3)
    Xp = torch.from_numpy(rs.uniform(-0.5, 0.5, size=(257, 6)).astype(np.float32))
    P = O.init_params(51)
    want = O.travel_times(P, Xp, m.B.cpu())
    m.network.precision = "fp32"
    got = m.TravelTimes(Xp.to(DEV)).cpu()
    torch.testing.assert_close(got, want, rtol=2e-5, atol=1e-7)

    fx = load_golden("plot_slices.npz")
    P = O.init_params(int(fx["seed"]))
    for prec, tol in (("fp32", 2e-5), ("tc", 1e-2)):
        mm = Model_multi(str(tmp_path), "", 3, 2, device=DEV)
        mm.init_network()
        _load_params(mm.network, P)
        mm.network.precision = prec
        mm.B = torch.from_numpy(fx["multi_B"]).to(DEV)
        X, Y, TT, V, TAU = mm.field_slice()
        np.testing.assert_array_equal(X, fx["multi_X"])
        np.testing.assert_array_equal(Y, fx["multi_Y"])
        # pointwise; TT = |x1-x0| / tau is 0/1 at the source point itself (-0.25, -0.25) in both implementations
        np.testing.assert_allclose(TAU, fx["multi_TAU"], rtol=tol / 10 if prec == "tc" else tol, atol=1e-7)
        np.testing.assert_allclose(TT, fx["multi_TT"], rtol=tol / 10 if prec == "tc" else tol, atol=1e-6)
        np.testing.assert_allclose(V, fx["multi_V"], rtol=tol, atol=1e-6)
        ms = Model_single(str(tmp_path), "", 6, device=DEV)
        ms.init_network(B=torch.from_numpy(fx["single_B"]))
        _load_params(ms.network, P)
        ms.network.precision = prec
        X, Y, TT, V, TAU = ms.field_slice()
        np.testing.assert_array_equal(X, fx["single_X"])         # S:1286-1290: [-2, 2)^2 in steps of 0.05
        np.testing.assert_allclose(TAU, fx["single_TAU"], rtol=tol / 10 if prec == "tc" else tol, atol=1e-7)
        np.testing.assert_allclose(TT, fx["single_TT"], rtol=tol / 10 if prec == "tc" else tol, atol=1e-6)
        np.testing.assert_allclose(V, fx["single_V"], rtol=tol, atol=1e-6)
    # plot() is headless: the same arrays land in an .npz next to the checkpoints
    mm.Params["ModelPath"] = str(tmp_path)
    mm.plot(3, 0.125, 0.5)
    z = np.load(str(tmp_path / "plots3_0.5_0.125_0.npz"))
    assert z["TT"].shape == (80, 80) and np.isfinite(z["V"]).all()


def test_autograd_and_optimizer_semantics_of_the_drop_in(capi):
    """Loss scaling, accumulation over two backward calls, p.grad aliasing the optimiser's buffer, zero_grad, and
    Loss() under no_grad (forward-only kernels) -- what a caller of torch.optim.AdamW relies on (M:1070-1080)."""
    from pntf import synth
    from pntf.model import Model_multi
    torch.manual_seed(0)
    d, E, N = 3, 2, 300
    m = Model_multi("", "", d, E, device=DEV)
    m.init_network()
    m.network.precision = "fp32"
    B = synth.sample_B(d, DEV, seed=1, E=E)
    c, s = synth.sample_pairs(E * N, d, DEV, seed=2)
    pts, spd = c.view(E, N, 2 * d), s.view(E, N, 2)
    fp = m.network.flat()
    named = dict(m.network.named_parameters())

    loss, loss_n, diff = m.Loss(pts, spd, B, 2.0, 1e-3)
    assert diff.shape == (E, N) and abs(float(loss) - 2.0 * float(loss_n)) < 1e-6 * float(loss)
    loss.backward()
    g1 = fp.grad.clone()
    assert float(g1.abs().max()) > 0
    tiny = 2e-6 * float(g1.abs().max())       # the gradient atomics' order differs between calls: last bits
    w = named["generator.1.weight"]
    assert w.grad is not None and w.grad.data_ptr() == fp.grad_views["generator.1.weight"].data_ptr()
    assert named["encoder1.0.weight"].grad is None and named["encoder1.0.bias"].grad is None
    # a second call accumulates, with the incoming scale
    loss2, _, _ = m.Loss(pts, spd, B, 2.0, 1e-3)
    (3.0 * loss2).backward()
    torch.testing.assert_close(fp.grad, 4.0 * g1, rtol=2e-5, atol=4 * tiny)
    torch.testing.assert_close(w.grad, 4.0 * g1[fp.offsets["generator.1.weight"]:][:w.numel()].view_as(w), rtol=2e-5, atol=4 * tiny)
    # clipping p.grad in place is what the optimiser applies
    torch.nn.utils.clip_grad_norm_([p for k, p in named.items() if p.grad is not None], 1e-3)
    assert abs(float(torch.linalg.norm(fp.grad)) - 1e-3) < 1e-5
    m.optimizer.zero_grad()
    assert float(fp.grad.abs().max()) == 0.0 and float(w.grad.abs().max()) == 0.0
    # gradient through loss_n alone = gradient through loss / beta
    _, ln, _ = m.Loss(pts, spd, B, 2.0, 1e-3)
    ln.backward()
    torch.testing.assert_close(fp.grad, g1 / 2.0, rtol=2e-5, atol=tiny)
    m.optimizer.zero_grad()
    # no_grad: same numbers, no gradient work
    with torch.no_grad():
        l3, ln3, d3 = m.Loss(pts, spd, B, 2.0, 1e-3)
    assert float(fp.grad.abs().max()) == 0.0 and not l3.requires_grad
    torch.testing.assert_close(l3, loss.detach(), rtol=1e-6, atol=0)
    torch.testing.assert_close(d3, diff, rtol=1e-6, atol=1e-9)
    # one optimiser step on the accumulated gradient equals torch.optim.AdamW on the same gradient
    before = fp.flat.clone()
    fp.grad.copy_(g1)
    m.optimizer.step()
    ref_p = torch.nn.Parameter(before.clone())
    ref_p.grad = g1.clone()
    ref_opt = torch.optim.AdamW([ref_p], lr=1e-3, weight_decay=0.1)
    ref_opt.step()
    used = fp.mask.bool()
    torch.testing.assert_close(fp.flat[used], ref_p.detach()[used], rtol=1e-6, atol=1e-9)
    torch.testing.assert_close(fp.flat[~used], before[~used], rtol=0, atol=0)


@pytest.mark.parametrize("d,eta,thr,cap", [(3, 0.03, 0.06, 60), (6, 0.015, 0.03, 40)])
@pytest.mark.parametrize("prec_name", ["tc", "fp32"])
def test_compact_plan_writes_the_same_waypoints_without_the_tail(capi, d, eta, thr, cap, prec_name):
    from oracle import pntf_oracle as O
    from pntf import synth
    prec = capi.PREC_TC if prec_name == "tc" else capi.PREC_FP32
    P = O.init_params(0)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    B = synth.sample_B(d, DEV, seed=1, E=1).contiguous()
    q = synth.sample_queries(777, d, DEV)
    q[5, d:] = q[5, :d]                               # a query whose endpoints already meet
    way, nst = capi.plan(flat, B, q, d, eta, thr, cap, prec=prec)
    torch.cuda.synchronize()
    from pntf.capi import workspace_bytes, workspace, lib, _ptr, _stream, OP_PLAN, _check
    sentinel = torch.full_like(way, -7.0)
    nst2 = torch.zeros_like(nst)
    ws = workspace(workspace_bytes(OP_PLAN, 1, q.shape[0], d, prec), q.device)
    _check(lib().pntf_plan_compact(_ptr(flat), _ptr(B), _ptr(q), q.shape[0], d, eta, thr, cap, 0, _ptr(sentinel),
                                   _ptr(nst2, torch.int32), prec, ws.data_ptr(), ws.numel(), _stream()))
    torch.cuda.synchronize()
    assert torch.equal(nst, nst2) and int(nst[5]) == 0
    rows = torch.arange(cap + 2, device=DEV).view(1, -1, 1)
    live = rows <= nst.view(-1, 1, 1)
    assert torch.equal(torch.where(live, sentinel, torch.zeros_like(way)), torch.where(live, way, torch.zeros_like(way)))
    assert bool((sentinel[~live.expand_as(way)] == -7.0).all())       # the tail was not touched


def test_query_sharding_is_a_no_op_without_a_process_group(capi):
    from pntf import synth
    from pntf.model import Model_multi
    torch.manual_seed(0)
    m = Model_multi("", "", 3, 1, device=DEV)
    m.init_network()
    m.B = synth.sample_B(3, DEV, seed=1, E=1)[0]
    q = synth.sample_queries(100, 3, DEV)
    a = m.TravelTimes(q)
    b = m.TravelTimes(q, shard=True, gather=True)
    assert torch.equal(a, b)
    w1, n1 = m.plan(q, step=0.03, thr=0.06, max_iter=30)
    w2, n2 = m.plan(q, step=0.03, thr=0.06, max_iter=30, shard=True, fill_tail=False)
    assert torch.equal(n1, n2)
    for i in range(0, 100, 17):
        assert torch.equal(m.path(w1, n1, i, 3), m.path(w2, n2, i, 3))


@pytest.mark.parametrize("prec", ["fp32", "tc"])
def test_graph_replayed_passes_equal_the_launch_by_launch_loop(capi, prec):
    """from the second pass on, Model.train's pass (index build, zero, loss + backward, AdamW, loss
    accumulation, x steps) is one CUDA-graph replay with alpha / beta / gamma / lr / step count read from device memory
    (pntf_loss_bwd_gather_sched, pntf_adamw_sched).  Same numbers as the launch-by-launch loop up to the order of the
    gradient atomics."""
    from pntf.model import Model_multi
    from pntf.trainer import run_training
    rows, d = 30000, 3
    envs = [train_fixture_env(80 + i, rows, d) for i in range(2)]
    datas = [torch.from_numpy(np.concatenate([p.astype(np.float16).astype(np.float32), s], 1)) for p, s, _ in envs]
    Bs = [torch.from_numpy(b) for _, _, b in envs]
    hists, finals = [], []
    for graph in (False, True):
        torch.manual_seed(5)
        m = Model_multi("", "", d, 2, device=DEV)
        m.init_network()
        m.network.precision = prec

        class Hooks:
            def __init__(self):
                self.gen = torch.Generator().manual_seed(3)
                self.steps = []

            def randperm(self, n):
                return torch.randperm(n, generator=self.gen)

            def rollback_index(self):
                return 1

            def on_step(self, rec):
                self.steps.append((rec["epoch"], rec["step"], rec["alpha"], rec["beta"], rec["lr"],
                                   [float(x) for x in rec["loss_out"].tolist()]))

            def on_epoch(self, rec):
                pass

        hooks = Hooks()
        hist = run_training(m, datas, Bs, epochs=5, log=lambda s: None, hooks=hooks, graph=graph)
        hists.append((hist, hooks.steps, m.optimizer.step_count))
        finals.append(m.network.flat().flat.clone())
    (h0, s0, c0), (h1, s1, c1) = hists
    assert c0 == c1 == 5 * 3 and len(s0) == len(s1) == 15
    tol = 2e-4 if prec == "fp32" else 2e-2
    for a, b in zip(s0, s1):
        assert a[:3] == b[:3] and a[4] == b[4]
        assert abs(a[3] - b[3]) <= tol * abs(a[3])
        np.testing.assert_allclose(b[5], a[5], rtol=tol)
    np.testing.assert_allclose([x[1] for x in h1], [x[1] for x in h0], rtol=tol)
    rel = float((finals[1] - finals[0]).abs().max() / finals[0].abs().max())
    assert rel < (5e-3 if prec == "fp32" else 5e-2), rel


def test_rollback_inside_graph_replayed_training(capi):
    """The same rollback rule when passes are CUDA-graph replays: the optimiser's step count lives in device memory the
    host rewrites before every replay, so a restored snapshot (weights, moments, step count) is what the next replay
    continues from."""
    from pntf import trainer as T
    from pntf.model import Model_multi
    rows, d = 10000, 3
    envs = [train_fixture_env(90 + i, rows, d) for i in range(2)]
    datas = [torch.from_numpy(np.concatenate([p.astype(np.float16).astype(np.float32), s], 1)) for p, s, _ in envs]
    Bs = [torch.from_numpy(b) for _, _, b in envs]
    torch.manual_seed(2)
    model = Model_multi("", "", d, 2, device=DEV)
    model.init_network()
    fp, opt = model.network.flat(), model.optimizer
    states = []
    orig_run = T._PassGraph.run

    def run(self, *a, **kw):
        c = len(states)
        states.append((fp.flat.clone(), opt.exp_avg.clone(), opt.exp_avg_sq.clone(), opt.step_count))
        out = orig_run(self, *a, **kw)
        if c == 2:                      # third graph pass = epoch 4, first attempt (epoch 1 runs launch by launch)
            out = [[50.0 * x for x in row] for row in out]
        return out

    class Hooks(T.TrainHooks):
        def rollback_index(self):
            return 3                    # ring at epoch 4 = [S0, S0, S1, S2]: the start of epoch 3

    T._PassGraph.run = run
    try:
        logs = []
        T.run_training(model, datas, Bs, epochs=5, log=logs.append, hooks=Hooks(0, DEV), graph=True)
    finally:
        T._PassGraph.run = orig_run
    # graph passes: epoch 2, epoch 3, epoch 4 (spiked), epoch 4 again, epoch 5
    assert len(states) == 5 and sum("RepeatEpoch" in line for line in logs) == 1
    assert all(torch.equal(x, y) for x, y in zip(states[3][:3], states[1][:3])) and states[3][3] == states[1][3] == 2
    assert not torch.equal(states[3][0], states[2][0])
    assert states[4][3] == 3 and opt.step_count == 4
