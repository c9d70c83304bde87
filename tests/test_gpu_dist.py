"""Two NCCL ranks on two GPUs (skipped on a one-GPU box): the data-parallel training step equals the single-rank step of
the concatenated batch, replicas stay bit-identical through Model.train's loop although every rank initialised from its
own random stream, and sharded planning / inference return the unsharded results."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, ws, port, outdir):
    import torch.distributed as dist
    from pntf import capi, synth
    from pntf.dist import GradAllReduce, shard_range
    from pntf.model import Model_multi
    from pntf.trainer import run_training
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    try:
        res = {}
        d, E, N = 3, 2, 5001                      # an odd batch: the ranks' shards differ in size
        # ---- 1. one data-parallel step == the single-rank step of the whole batch (fp32 mode: 1e-5) ----
        torch.manual_seed(7)                      # same weights on both ranks for this part
        m = Model_multi("", "", d, E, device=str(dev))
        m.init_network()
        m.network.precision = "fp32"
        fp = m.network.flat()
        B = synth.sample_B(d, dev, seed=1, E=E).contiguous()
        c, s = synth.sample_pairs(E * N, d, dev, seed=2)
        table = torch.cat([c, s], dim=1).contiguous()
        b, e = shard_range(N, rank, ws)
        n = e - b
        index = torch.cat([torch.arange(b, e, device=dev) + env * N for env in range(E)]).to(torch.int32)
        w = n * ws / float(N)
        beta = 1.7
        fp.grad.zero_()
        m._loss_step_gather(table, index, B, E, n, 1.0, beta * w, 1e-3)
        fp.loss_slot[1:].mul_(w)
        red = GradAllReduce(capi.PARAM_COUNT, dev)
        grad, loss = red.reduce_flat(fp)
        corr = (ws - 1) * 0.01 * float((B.double() ** 2).sum()) / (E * N)
        res["dp_grad"] = (grad / ws).cpu().clone()
        res["dp_loss"] = (loss.cpu().double() - torch.tensor([beta * corr, corr], dtype=torch.float64))
        if rank == 0:
            g = torch.zeros_like(fp.flat)
            lo, _ = capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, E, N, d, 1.0, beta, 1e-3, grad_out=g, prec=capi.PREC_FP32)
            res["full_grad"] = g.cpu()
            res["full_loss"] = lo.cpu().double()

        # ---- 2. Model.train's loop: every rank draws its own initial weights; replicas must end identical ----
        torch.manual_seed(100 + rank)
        m2 = Model_multi("", "", d, 2, device=str(dev))
        m2.init_network()
        res["init_head"] = m2.network.flat().flat[:8].cpu().clone()
        rows = 20000                               # two steps of 2 x 10 000 pairs per epoch
        datas, Bs = [], []
        for env in range(2):
            cc, ss = synth.sample_pairs(rows, d, dev, seed=30 + env)
            datas.append(torch.cat([cc, ss], dim=1))
            Bs.append(synth.sample_B(d, dev, seed=40 + env, E=1)[0])
        hist = run_training(m2, datas, Bs, epochs=3, log=lambda s: None)
        res["flat"] = m2.network.flat().flat.cpu().clone()
        res["exp_avg"] = m2.optimizer.exp_avg.cpu().clone()
        res["steps"] = m2.optimizer.step_count
        res["hist"] = hist

        # ---- 3. sharded planning and inference ----
        torch.manual_seed(7)
        m3 = Model_multi("", "", d, 1, device=str(dev))
        m3.init_network()
        m3.B = synth.sample_B(d, dev, seed=1, E=1)[0]
        q = synth.sample_queries(1001, d, dev)
        way, nst = m3.plan(q, step=0.03, thr=0.06, max_iter=40, shard=True, gather=True)
        tt = m3.TravelTimes(q, shard=True, gather=True)
        local_tt = m3.TravelTimes(q, shard=True)
        res["way"], res["nst"], res["tt"], res["local_rows"] = way.cpu(), nst.cpu(), tt.cpu(), local_tt.shape[0]
        if rank == 0:
            w0, n0 = m3.plan(q, step=0.03, thr=0.06, max_iter=40)
            res["way0"], res["nst0"], res["tt0"] = w0.cpu(), n0.cpu(), m3.TravelTimes(q).cpu()
        torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def test_two_rank_training_planning_and_inference(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(str(tmp_path / "r0.pt"), weights_only=False)
    r1 = torch.load(str(tmp_path / "r1.pt"), weights_only=False)
    # 1. all-reduced gradient and loss == single-rank gradient and loss of the concatenated batch
    g, f = r0["dp_grad"].double(), r0["full_grad"].double()
    assert float((g - f).abs().max() / f.abs().max()) < 1e-5
    assert torch.equal(r0["dp_grad"], r1["dp_grad"])
    np.testing.assert_allclose(r0["dp_loss"].numpy(), r0["full_loss"].numpy(), rtol=1e-5)
    # 2. different initial draws, identical replicas after three epochs (6 steps)
    assert not torch.equal(r0["init_head"], r1["init_head"])
    assert torch.equal(r0["flat"], r1["flat"]) and torch.equal(r0["exp_avg"], r1["exp_avg"])
    assert r0["steps"] == r1["steps"] == 3 * 2
    assert r0["hist"] == r1["hist"] and r0["hist"][-1][1] < r0["hist"][0][1]
    # 3. sharded results == unsharded results, on both ranks
    for r in (r0, r1):
        assert torch.equal(r["nst"], r0["nst0"]) and torch.equal(r["tt"], r0["tt0"])
        assert torch.equal(r["way"], r0["way0"])
    assert r0["local_rows"] == 501 and r1["local_rows"] == 500
