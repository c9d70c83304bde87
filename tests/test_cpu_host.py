"""Host-side rows without a GPU: trainer schedule (T2) against the recorded reference run, dataset readers (8f rank 2),
checkpoint interop in both directions (8f rank 4), the slice grid of Model.plot, and the data-parallel plumbing under
gloo with world size 2 (replica broadcast, in-place gradient all-reduce, unequal-shard weighting, query sharding)."""
import os
import pickle
import socket
import sys
import types

import numpy as np
import pytest
import torch

from helpers import GOLDEN, load_golden, train_fixture_env, write_multi_dataset, write_single_dataset

from pntf import data as pdata
from pntf.trainer import lr_schedule

REF = "/root/reference"


# ------------------------------------------------------------------------------------------- schedule
def test_lr_and_alpha_schedule_match_the_reference_formulas_and_the_recorded_run():
    # M:1011-1026 / S:1025-1039
    assert lr_schedule(1) == 1e-3 and lr_schedule(8000) == 1e-3
    assert abs(lr_schedule(8250) - 7.5e-4) < 1e-12 and lr_schedule(8500) == 5e-4 and lr_schedule(9999) == 5e-4
    fx = load_golden("train_multi_d3.npz")
    steps = fx["steps"]                       # beta, gamma, lr, loss, loss_n, first speed value of every reference step
    assert np.all(steps[:, 1] == 1e-3)                                 # gamma
    assert np.all(steps[:, 2] == [lr_schedule(e) for e in range(1, len(steps) + 1)])
    # beta of epoch e = 1 / total_diff of epoch e-1, total_diff = sum(loss_n) / (len(dataloader) * 5) (M:1093-1109)
    total_diff = steps[:, 4] / 5.0
    assert steps[0, 0] == 1.0
    np.testing.assert_allclose(steps[1:, 0], 1.0 / total_diff[:-1], rtol=2e-6)
    np.testing.assert_allclose(fx["epoch_log"][:, 1], total_diff, rtol=1e-4)     # the printed "Loss" is total_diff
    # alpha ramp (M:1011-1014): 0.5 until step turns positive after 2000 epochs
    step, alphas = -0.5, []
    for epoch in range(1, 4500):
        alphas.append(min(max(0.5, 0.5 + 0.5 * step), 1.07))
        step += 1.0 / 4000 / (int(epoch / 4000) + 1.)
    assert alphas[0] == 0.5 and alphas[2000] == 0.5                     # flat for the first 2000 epochs
    assert abs(alphas[3000] - 0.625) < 1e-9 and abs(alphas[3999] - (0.5 + 0.5 * (-0.5 + 3999 / 4000.))) < 1e-9
    assert abs((alphas[4400] - alphas[4399]) - 0.5 / 8000) < 1e-12      # half the slope from epoch 4000 on
    np.testing.assert_allclose(fx["epoch_log"][:, 2], alphas[:len(fx["epoch_log"])])
    # checkpoints and plots at epoch 1 and at the last epoch (M:1129-1132)
    files = [str(f) for f in fx["files"]]
    assert sum(f.startswith("Model_Epoch_00001_ValLoss_") for f in files) == 1
    assert sum(f.startswith("Model_Epoch_00008_ValLoss_") for f in files) == 1


# ------------------------------------------------------------------------------------------- readers
def test_multi_reader_follows_path_plus_index_and_rounds_through_fp16(tmp_path):
    envs = [train_fixture_env(100 + i, 50, 3) for i in range(12)]
    root = str(tmp_path / "env")                 # a prefix, as `path + str(index)` allows (models/data_multi.py:17)
    write_multi_dataset(root, envs)
    got = pdata.load_multi(root, 12)
    assert len(got) == 12
    for i, (table, B) in enumerate(got):         # numeric order, not the lexicographic 0, 1, 10, 11, 2 ...
        pts, spd, Bw = envs[i]
        assert table.shape == (50, 8) and table.dtype == torch.float32
        np.testing.assert_array_equal(table[:, :6].numpy(), pts.astype(np.float16).astype(np.float32))   # :19
        assert not np.array_equal(table[:, :6].numpy(), pts)
        np.testing.assert_array_equal(table[:, 6:].numpy(), spd)
        np.testing.assert_array_equal(B.numpy(), Bw)
    with pytest.raises(FileNotFoundError):
        pdata.load_multi(root, 13)


def test_single_reader_layout(tmp_path):
    pts, spd, _ = train_fixture_env(7, 33, 6)
    write_single_dataset(str(tmp_path), pts, spd)
    t = pdata.load_single(str(tmp_path))
    assert t.shape == (33, 14)
    np.testing.assert_array_equal(t[:, :12].numpy(), pts)       # no fp16 rounding in models/data_mlp.py:25-43
    np.testing.assert_array_equal(t[:, 12:].numpy(), spd)


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference checkout (build container)")
def test_readers_agree_with_the_reference_loaders(tmp_path):
    data_multi, data_mlp = _import_reference(("data_multi", "data_mlp"))
    envs = [train_fixture_env(200 + i, 40, 3) for i in range(3)]
    root = str(tmp_path / "m") + "/"
    write_multi_dataset(root, envs)
    ref = data_multi.Database(root, torch.device("cpu"), 3)
    for i, (table, B) in enumerate(pdata.load_multi(root, 3)):
        rdata, rB, ridx = ref[i]
        assert ridx == i
        torch.testing.assert_close(table, rdata, rtol=0, atol=0)
        torch.testing.assert_close(B, rB, rtol=0, atol=0)
    pts, spd, _ = train_fixture_env(9, 25, 6)
    sroot = str(tmp_path / "s")
    write_single_dataset(sroot, pts, spd)
    torch.testing.assert_close(pdata.load_single(sroot), data_mlp.Database(sroot).data, rtol=0, atol=0)


# ------------------------------------------------------------------------------------------- checkpoints
def _oracle_params(seed):
    from oracle import pntf_oracle as O
    return O.init_params(seed)


def test_reference_written_checkpoint_loads(tmp_path):
    from pntf.model import Model_multi
    path = os.path.join(GOLDEN, "ref_ckpt", "Model_Epoch_00003_ValLoss_1.250000e-02.pt")
    m = Model_multi(str(tmp_path), "", 3, 2, device="cpu")
    m.load(path)
    P = _oracle_params(51)
    sd = m.network.state_dict()
    assert list(sd.keys()) == list(P.keys())
    for k in P:
        torch.testing.assert_close(sd[k], P[k], rtol=0, atol=0)
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert set(ck.keys()) == {"epoch", "model_state_dict", "optimizer_state_dict", "B_state_dict", "train_loss", "val_loss"}
    assert ck["epoch"] == 3 and tuple(ck["B_state_dict"].shape) == (3, 128)


def _import_reference(names=("model_res_sigmoid", "model_res_sigmoid_multi")):
    """Modules of the reference's `models` package, imported in isolation from the repo's drop-in package of the same
    name (2-module shim: empty matplotlib / matplotlib.pylab, pickle5 -> pickle)."""
    import importlib
    for name in ("matplotlib", "matplotlib.pylab"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules.setdefault("pickle5", pickle)
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "models" or k.startswith("models.")}
    sys.path.insert(0, REF)
    try:
        mods = tuple(importlib.import_module("models." + n) for n in names)
    finally:
        sys.path.remove(REF)
        for k in list(sys.modules):
            if k == "models" or k.startswith("models."):
                del sys.modules[k]
        sys.modules.update(saved)
    return mods


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference checkout (build container)")
def test_checkpoint_round_trip_with_the_reference_in_both_directions(tmp_path):
    from pntf.model import Model_multi, Model_single
    S, M = _import_reference()
    # repo -> reference, multi
    torch.manual_seed(3)
    mine = Model_multi(str(tmp_path), "", 3, 2, device="cpu")
    mine.init_network()
    mine.B = 0.5 * torch.randn(3, 128)
    mine.total_train_loss = [0.5, 0.25]
    mine.save(epoch=12, val_loss=0.03125)
    f = str(tmp_path / "Model_Epoch_00012_ValLoss_3.125000e-02.pt")
    assert os.path.exists(f)
    ref = M.Model(str(tmp_path), "", 3, 2, device="cpu")
    ref.load(f)
    for (k, v), (k2, v2) in zip(ref.network.state_dict().items(), mine.network.state_dict().items()):
        assert k == k2
        torch.testing.assert_close(v, v2, rtol=0, atol=0)
    # repo -> reference, single (the reference's load also restores B, S:1154-1156)
    mine1 = Model_single(str(tmp_path), "", 6, device="cpu")
    mine1.init_network()
    mine1.save(epoch=1, val_loss=0.5)
    ref1 = S.Model(str(tmp_path), "", 6, device="cpu")
    ref1.load(str(tmp_path / "Model_Epoch_00001_ValLoss_5.000000e-01.pt"))
    torch.testing.assert_close(ref1.B, mine1.B, rtol=0, atol=0)
    torch.testing.assert_close(ref1.network.B, mine1.network.B.cpu(), rtol=0, atol=0)
    for (k, v), (k2, v2) in zip(ref1.network.state_dict().items(), mine1.network.state_dict().items()):
        assert k == k2
        torch.testing.assert_close(v, v2, rtol=0, atol=0)
    # reference -> repo, single, with optimiser state (one AdamW step taken)
    ref2 = S.Model(str(tmp_path), "", 6, device="cpu")
    ref2.B = 0.5 * torch.randn(128, 6)
    ref2.network = S.NN("cpu", 6, ref2.B)
    ref2.network.load_state_dict(_oracle_params(4), strict=True)
    ref2.optimizer = torch.optim.AdamW(ref2.network.parameters(), lr=1e-3, weight_decay=0.1)
    for k, p in ref2.network.named_parameters():
        if not k.startswith("encoder1.0"):
            p.grad = torch.full_like(p, 1e-3)
    ref2.optimizer.step()
    ref2.save(epoch=7, val_loss=torch.tensor(0.75))
    mine2 = Model_single(str(tmp_path), "", 6, device="cpu")
    mine2.load(str(tmp_path / "Model_Epoch_00007_ValLoss_7.500000e-01.pt"))
    torch.testing.assert_close(mine2.B, ref2.B, rtol=0, atol=0)
    for (k, v), (k2, v2) in zip(ref2.network.state_dict().items(), mine2.network.state_dict().items()):
        assert k == k2
        torch.testing.assert_close(v, v2, rtol=0, atol=0)


def test_fused_optimizer_state_dict_has_the_torch_adamw_shape():
    """The checkpoint's optimizer_state_dict must be loadable by torch.optim.AdamW.load_state_dict (the reference's
    rollback does exactly that, M:1106): param indices in parameter order, step / exp_avg / exp_avg_sq per entry."""
    from pntf.model import Model_multi
    m = Model_multi("", "", 3, 2, device="cpu")
    m.init_network()
    m.optimizer.step_count = 2
    m.optimizer.exp_avg.fill_(0.5)
    sd = m.optimizer.state_dict()
    probe = torch.optim.AdamW([torch.nn.Parameter(p.detach().clone()) for p in m.network.parameters()], lr=1e-3,
                              weight_decay=0.1)
    probe.load_state_dict(sd)
    names = [k for k, _ in m.network.named_parameters()]
    unused = [i for i, k in enumerate(names) if k.startswith("encoder1.0")]
    assert sorted(sd["state"].keys()) == [i for i in range(30) if i not in unused]
    assert float(sd["state"][0]["step"]) == 2.0 and sd["state"][0]["exp_avg"].shape == (128, 256)
    assert probe.param_groups[0]["lr"] == 1e-3 and probe.param_groups[0]["weight_decay"] == 0.1


# ------------------------------------------------------------------------------------------- slices
def test_slice_grid_matches_model_plot():
    fx = load_golden("plot_slices.npz")
    limit, n = 0.5, 80
    ax = np.arange(-limit, limit, 2 * limit / n)
    X, Y = np.meshgrid(ax, ax)
    np.testing.assert_array_equal(X, fx["multi_X"])           # M:1251-1255: spacing = limit / 40 -> 80 x 80
    np.testing.assert_array_equal(Y, fx["multi_Y"])


# ------------------------------------------------------------------------------------------- data-parallel plumbing
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dp_worker(rank, ws, port, out):
    import torch.distributed as dist
    from pntf.dist import GradAllReduce, broadcast_replica_state, gather_rows, shard_range, shard_rows
    from pntf.params import FlatParams
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        torch.manual_seed(100 + rank)                      # every rank initialises from its own stream ...
        fp = FlatParams("cpu")
        fp.init_weights()
        m, v = torch.rand_like(fp.flat), torch.rand_like(fp.flat)
        B = torch.randn(3, 128)

        class Opt:
            step_count = 5 + rank
        opt = Opt()
        broadcast_replica_state([fp.flat, m, v, B], opt)   # ... and becomes a replica of rank 0
        res = {"flat": fp.flat.clone(), "m": m.clone(), "B": B.clone(), "step": opt.step_count}

        # unequal shards (10 001 pairs over 2 ranks): weighted local means all-reduce to the global mean
        n_g = 10001
        rs = np.random.RandomState(5)
        per_pair = torch.from_numpy(rs.rand(n_g).astype(np.float32))         # a per-pair loss term
        gvec = torch.from_numpy(rs.rand(n_g, 7).astype(np.float32))          # a per-pair gradient contribution
        b, e = shard_range(n_g, rank, ws)
        n = e - b
        w = n * ws / float(n_g)
        fp.grad.zero_()
        fp.grad[:7] = gvec[b:e].mean(0) * w                # what the kernel writes with beta * w
        fp.loss_slot[0] = per_pair[b:e].mean() * w
        fp.loss_slot[1] = per_pair[b:e].mean() * w
        red = GradAllReduce(fp.flat.numel(), "cpu")
        grad, loss = red.reduce_flat(fp)
        assert grad.data_ptr() == fp.grad.data_ptr()       # in place, no packing copies
        res["grad"] = (grad[:7] / ws).clone()
        res["loss"] = loss.clone()
        res["want_grad"] = gvec.mean(0)
        res["want_loss"] = per_pair.mean()

        q = torch.arange(37 * 2, dtype=torch.float32).view(37, 2)
        local, lo, hi = shard_rows(q)
        res["rows"] = (lo, hi)
        res["gathered"] = gather_rows(local * 2, 37)
        out[rank] = res
    finally:
        dist.destroy_process_group()


def _dp_entry(rank, ws, port, path):
    out = {}
    _dp_worker(rank, ws, port, out)
    torch.save(out[rank], os.path.join(path, "r%d.pt" % rank))


def test_replica_broadcast_inplace_allreduce_and_sharding_gloo_world2(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_dp_entry, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(str(tmp_path / "r0.pt"), weights_only=False)
    r1 = torch.load(str(tmp_path / "r1.pt"), weights_only=False)
    for k in ("flat", "m", "B"):
        torch.testing.assert_close(r0[k], r1[k], rtol=0, atol=0)
    assert r0["step"] == r1["step"] == 5
    for r in (r0, r1):
        torch.testing.assert_close(r["grad"], r["want_grad"], rtol=1e-5, atol=1e-7)
        torch.testing.assert_close(r["loss"], torch.stack([r["want_loss"], r["want_loss"]]), rtol=1e-5, atol=1e-7)
        torch.testing.assert_close(r["gathered"], torch.arange(74, dtype=torch.float32).view(37, 2) * 2)
    assert r0["rows"] == (0, 19) and r1["rows"] == (19, 37)
