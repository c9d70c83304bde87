"""Trainer control flow of Model.train() (models/model_res_sigmoid.py:938-1137 single,
models/model_res_sigmoid_multi.py:953-1141 multi) on the fused kernels.

Kept from the reference: alpha ramp (0.5 -> cap), beta = 1/mean-diff of the previous epoch,
gamma = 1e-3, lr clip schedule, 6 steps of 10 000 pairs (per environment) per epoch, the
5-deep (network, optimiser) snapshot ring and the "loss ratio not in (0, 1.2) -> reload a random
snapshot and redo the epoch" rule, the `/ (len(dataloader) * 5)` normalisation of the multi trainer
(M:1093-1094, with 6 steps) and `/ len(dataloader)` of the single one (S:1083-1084).  Random draws
are consumed in the reference's order -- multi: environment order of the epoch (the DataLoader's
shuffle, M:967-971), then one row permutation per environment of a batch (M:1053-1058); single: one
row permutation per epoch, applied on top of the previous one (FastTensorDataLoader shuffles in place,
S:61-64) -- so a recorded reference run can be replayed step by step (`TrainHooks`,
tests/test_gpu_trainer.py).  Snapshots are device-to-device copies of the flat buffers.

Under torch.distributed every rank processes its slice of each step's pairs and the flat gradient
is all-reduced; rank 0's initial weights, Fourier matrices and optimiser state are broadcast first,
every rank draws the same permutations (same seed), and the rollback decision uses the all-reduced
loss and a shared random stream, so the replicas stay identical.

Device-side epoch: one pass over a batch of environments -- its up-to-six steps of
index build, zero, loss + backward, all-reduce, AdamW and loss accumulation -- is captured ONCE as a CUDA graph
(`_PassGraph`) and replayed for every later pass; alpha / beta / gamma, the learning rate, the optimiser's step count,
the row permutations, the environments' table offsets and Fourier matrices live in device buffers the host refreshes
before a replay (pntf_loss_bwd_gather_sched / pntf_adamw_sched read them at run time), and the pass's losses come back
in one read.  `graph=False` (or PNTF_TRAIN_GRAPH=0) keeps the launch-by-launch loop; both produce the same numbers.

Data feeding: all environments' tables are stacked once into one
GPU-resident [rows, 2d+2] table; a step's batch is an int32 index vector (`perm[lo:hi] + row offset` per
environment) handed to pntf_loss_bwd_gather, so the reference's per-step `data[perm[lo:hi]]` gather, `stack`
and `contiguous` copies (M:1053-1066) never materialise.
"""
import os
import random

import numpy as np
import torch

from . import capi
from .dist import GradAllReduce, broadcast_replica_state, shard_range, world
from .nn import _PREC


class TrainHooks:
    """Source of the trainer's random draws and observer of its schedule.  The default draws from a seeded
    generator; tests substitute recorded draws of a reference run and collect the per-step / per-epoch records."""

    def __init__(self, seed=0, device="cpu"):
        self.rng = random.Random(seed)                       # identical on every rank
        self.gen = torch.Generator(device=device)
        self.gen.manual_seed(seed)
        self.device = device

    def randperm(self, n):
        return torch.randperm(n, device=self.device, generator=self.gen)

    def rollback_index(self):
        return self.rng.randint(0, 4)                        # S:1093 / M:1104

    def on_step(self, rec):
        pass

    def on_epoch(self, rec):
        pass


def lr_schedule(epoch):
    """S:1039 / M:1025."""
    return float(np.clip(1e-3 * (1 - (epoch - 8000) / 1000.), a_min=5e-4, a_max=1e-3))


class _PassGraph:
    """One pass (<= 6 steps) over a batch of E environments as a replayable CUDA graph."""

    def __init__(self, model, table, E, rows, bounds, rank, ws):
        import torch.distributed as dist
        net, opt = model.network, model.optimizer
        fp = net.flat()
        dev = fp.device
        d = model.dim
        self.fp, self.opt, self.E, self.steps, self.ws = fp, opt, E, len(bounds), ws
        self.perm = torch.zeros(E, rows, dtype=torch.int32, device=dev)
        self.off = torch.zeros(E, 1, dtype=torch.int32, device=dev)
        self.off_host = torch.zeros(E, 1, dtype=torch.int32).pin_memory()
        self.B = torch.zeros(E, d, 128, dtype=torch.float32, device=dev)
        # per-step {alpha, beta * w, gamma}, per-step loss correction {beta c, c}, then {lr, step count}
        n_cfg = 5 * self.steps + 2
        self.cfg = torch.zeros(n_cfg, dtype=torch.float32, device=dev)
        self.cfg_host = torch.zeros(n_cfg, dtype=torch.float32).pin_memory()
        self.sched = self.cfg[:3 * self.steps].view(self.steps, 3)
        self.corr = self.cfg[3 * self.steps:5 * self.steps].view(self.steps, 2)
        self.lr_step = self.cfg[5 * self.steps:]
        self.losses = torch.zeros(self.steps, 2, dtype=torch.float32, device=dev)
        self.bounds = bounds                                  # (lo, b, e, n_g) per step: rows lo+b .. lo+e of a permutation
        self.index = [torch.zeros(E * (e - b), dtype=torch.int32, device=dev) for (lo, b, e, n_g) in bounds]
        self.diff = torch.zeros(E * max(e - b for (lo, b, e, n_g) in bounds), dtype=torch.float32, device=dev)
        prec = _PREC[net.precision]
        budget = net.stash_budget

        def body():
            for ii, (lo, b, e, n_g) in enumerate(bounds):
                n = e - b
                w = n * ws / float(n_g)
                for j in range(E):
                    torch.add(self.perm[j, lo + b:lo + e], self.off[j], out=self.index[ii][j * n:(j + 1) * n])
                fp.grad.zero_()
                capi.loss_gather_sched(model.variant, fp.flat, self.B, table, self.index[ii], E, n, d, self.sched[ii],
                                       fp.grad, self.diff[:E * n], fp.loss_slot, prec=prec, budget=budget)
                if w != 1.0:
                    fp.loss_slot[1:].mul_(w)
                if ws > 1:
                    dist.all_reduce(fp.gradbuf, op=dist.ReduceOp.SUM)
                self.lr_step[1:].add_(1.0)
                opt.step_sched(self.lr_step, grad_scale=1.0 / ws)
                torch.sub(fp.loss_slot / ws if ws > 1 else fp.loss_slot, self.corr[ii], out=self.losses[ii])

        # every buffer the body touches exists (the first pass ran launch by launch: workspace, NCCL); capture
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            body()

    def run(self, perms, offsets, B, alpha, beta, gamma, lr, b_term):
        """Refreshes the device-side inputs, replays, returns the per-step [loss, loss_n] (one device read)."""
        E, ws = self.E, self.ws
        for j in range(E):
            self.perm[j].copy_(perms[j])
            self.off_host[j, 0] = int(offsets[j])
        self.off.copy_(self.off_host, non_blocking=True)
        self.B.copy_(B)
        h = self.cfg_host
        for ii, (lo, b, e, n_g) in enumerate(self.bounds):
            w = (e - b) * ws / float(n_g)
            h[3 * ii + 0], h[3 * ii + 1], h[3 * ii + 2] = alpha, beta * w, gamma
            c = (ws - 1) * b_term / (E * n_g)
            h[3 * self.steps + 2 * ii], h[3 * self.steps + 2 * ii + 1] = beta * c, c
        h[5 * self.steps], h[5 * self.steps + 1] = lr, float(self.opt.step_count)
        self.cfg.copy_(h, non_blocking=True)
        self.graph.replay()
        self.opt.step_count += self.steps
        return self.losses.tolist()


def run_training(model, datas, Bs, epochs=None, log=print, inner_batch=10000, steps_per_epoch=6, seed=0, hooks=None,
                 graph=None):
    multi = model.variant == capi.LOSS_MULTI
    net, opt = model.network, model.optimizer
    fp = net.flat()
    dev = fp.device
    rank, ws = world()
    n_epochs = int(epochs if epochs is not None else model.Params['Training']['Number of Epochs'])
    if hooks is None:
        hooks = TrainHooks(seed, dev)
    reducer = GradAllReduce(capi.PARAM_COUNT, dev)
    Bs = [b.to(device=dev, dtype=torch.float32).contiguous() for b in Bs]
    if ws > 1:
        # replicas must start from the same weights / Fourier matrices / optimiser state (each rank drew its own)
        broadcast_replica_state([fp.flat, opt.exp_avg, opt.exp_avg_sq] + Bs, opt)
        if not multi:
            net.B = Bs[0]                        # [d, 128] as NN_single stores it (S:139)
            model.B = Bs[0].T

    n_env = len(datas)
    env_batch = int(model.Params['Training']['Batch Size']) if multi else 1
    rows = int(datas[0].shape[0])                # M:997: rows of the first environment
    if multi:
        n_loader = -(-n_env // env_batch)
        norm = n_loader * 5.0                    # M:1093-1094 divides by len(dataloader)*5 with 6 steps
        steps = min(int(rows / inner_batch), steps_per_epoch)
    else:
        inner_batch = int(model.Params['Training']['Batch Size'])
        n_loader = -(-rows // inner_batch)
        norm = float(n_loader)                   # S:1083-1084
        steps = min(n_loader, steps_per_epoch)
    if steps < 1:
        raise RuntimeError("the dataset has %d rows per environment, fewer than one batch of %d" % (rows, inner_batch))
    if inner_batch < ws:
        raise RuntimeError("a batch of %d pairs cannot be split over %d ranks" % (inner_batch, ws))

    # one GPU-resident table: the environments' rows stacked (row offset per environment)
    offs = [0]
    for t in datas:
        offs.append(offs[-1] + int(t.shape[0]))
    table = torch.cat([t.to(device=dev, dtype=torch.float32) for t in datas]).contiguous()
    single_order = None                          # single: the loader's cumulative in-place shuffle (S:61-64)

    if graph is None:
        graph = os.environ.get("PNTF_TRAIN_GRAPH", "1") != "0"
    pass_graphs = {}                             # environments per batch -> _PassGraph (built after the first eager pass)
    eager_passes = 0
    bounds = []
    for ii in range(steps):
        lo, hi = ii * inner_batch, min((ii + 1) * inner_batch, rows)
        b, e_ = shard_range(hi - lo, rank, ws)
        bounds.append((lo, b, e_, hi - lo))
    if any(e_ - b == 0 for (lo, b, e_, n_g) in bounds):
        graph = False                            # a rank with an empty shard skips its kernels: launch by launch only

    beta, current_diff, step = 1.0, 1.0, -2000.0 / 4000.0
    snap = lambda: (fp.flat.clone(), opt.snapshot())
    current_state = snap()
    queue = []
    history = []
    for epoch in range(1, n_epochs + 1):
        alpha = min(max(0.5, 0.5 + 0.5 * step), model.alpha_cap)
        step += 1.0 / 4000 / (int(epoch / 4000) + 1.)
        gamma = 0.001
        queue.append(current_state)
        if len(queue) > 5:
            queue.pop(0)
        current_state = snap()
        opt.param_groups[0]['lr'] = lr_schedule(epoch)
        prev_diff = current_diff
        repeats = 0
        while True:
            totals = torch.zeros(2, dtype=torch.float32, device=dev)
            if multi:
                env_order = [int(e) for e in hooks.randperm(n_env).tolist()]
                batches = [env_order[i:i + env_batch] for i in range(0, n_env, env_batch)]
            else:
                batches = [[0]]
            last_B = None
            for grp in batches:
                E = len(grp)
                if multi:
                    perms = [hooks.randperm(rows).to(dev) for _ in grp]
                else:
                    r = hooks.randperm(rows).to(dev)
                    single_order = r if single_order is None else single_order[r]
                    perms = [single_order]
                B = torch.stack([Bs[e] for e in grp]).contiguous()
                last_B = Bs[grp[0]]              # M:1069: self.B = B[0, :]
                # 0.01 |B|^2 of the multi loss (M:947) enters every rank's loss_n once: taken out ws-1 times below
                b_term = 0.01 * float((B.double() ** 2).sum().item()) if (multi and ws > 1) else 0.0
                if graph and eager_passes >= 1:
                    pg = pass_graphs.get(E)
                    if pg is None:
                        pg = pass_graphs[E] = _PassGraph(model, table, E, rows, bounds, rank, ws)
                    lr_now = opt.param_groups[0]['lr']
                    step_losses = pg.run(perms, [offs[e] for e in grp], B, alpha, beta, gamma, lr_now, b_term)
                    for ii, lo_ in enumerate(step_losses):
                        lt = torch.tensor(lo_, dtype=torch.float32, device=dev)
                        totals += lt
                        hooks.on_step({"epoch": epoch, "step": ii, "envs": list(grp), "alpha": alpha, "beta": beta,
                                       "gamma": gamma, "lr": lr_now, "loss_out": lt, "index": None})
                    continue
                eager_passes += 1
                for ii in range(steps):
                    lo, hi = ii * inner_batch, min((ii + 1) * inner_batch, rows)
                    n_g = hi - lo
                    b, e_ = shard_range(n_g, rank, ws)
                    n = e_ - b
                    # a rank's loss and gradient are means over ITS n pairs; weighting them by n ws / n_g makes the
                    # all-reduced sum / ws the mean over the step's n_g pairs whatever the shard sizes (w = 1 when equal)
                    w = n * ws / float(n_g)
                    fp.grad.zero_()
                    if n > 0:
                        index = torch.cat([perms[j][lo + b:lo + e_] + offs[grp[j]] for j in range(E)]).to(torch.int32)
                        model._loss_step_gather(table, index, B, E, n, alpha, beta * w, gamma)
                        if w != 1.0:
                            fp.loss_slot[1:].mul_(w)
                    else:
                        index = None
                        fp.loss_slot.zero_()
                    _, loss_out = reducer.reduce_flat(fp)
                    if b_term != 0.0:
                        corr = (ws - 1) * b_term / (E * n_g)
                        loss_out = loss_out - torch.tensor([beta * corr, corr], dtype=torch.float32, device=dev)
                    opt.step(grad_scale=1.0 / ws)
                    totals += loss_out
                    hooks.on_step({"epoch": epoch, "step": ii, "envs": list(grp), "alpha": alpha, "beta": beta,
                                   "gamma": gamma, "lr": opt.param_groups[0]['lr'], "loss_out": loss_out, "index": index})
            totals /= norm
            total_train_loss, total_diff = totals[0].item(), totals[1].item()
            current_diff = total_diff
            ratio = current_diff / prev_diff
            if 0 < ratio < 1.2:
                break
            repeats += 1
            k = hooks.rollback_index()
            k = min(k, len(queue) - 1)        # the reference indexes past the ring before 5 epochs exist
            fp.flat.copy_(queue[k][0])
            opt.restore(queue[k][1])
            log("RepeatEpoch = {} -- Loss = {:.4e} -- Alpha = {:.4e}".format(epoch, total_diff, alpha))
        model.total_train_loss.append(total_train_loss)
        hooks.on_epoch({"epoch": epoch, "alpha": alpha, "beta": beta, "gamma": gamma, "lr": opt.param_groups[0]['lr'],
                        "total_diff": total_diff, "total_train_loss": total_train_loss, "repeats": repeats})
        beta = 1.0 / total_diff
        history.append((epoch, total_diff, alpha))
        if rank == 0 and epoch % model.Params['Training']['Print Every * Epoch'] == 0:
            log("Epoch = {} -- Loss = {:.4e} -- Alpha = {:.4e}".format(epoch, total_diff, alpha))
        if multi:
            model.B = last_B
        if rank == 0 and model.Params['ModelPath'] and (
                epoch % model.Params['Training']['Save Every * Epoch'] == 0 or epoch == n_epochs or epoch == 1):
            try:
                model.plot(epoch, total_diff, alpha)
                model.save(epoch=epoch, val_loss=total_diff)
            except OSError as exc:   # unwritable ModelPath: keep training, report
                log("checkpoint not written: %s" % exc)
    return history
