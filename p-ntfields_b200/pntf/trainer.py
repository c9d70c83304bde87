"""Trainer control flow of Model.train() (models/model_res_sigmoid.py:938-1137 single,
models/model_res_sigmoid_multi.py:953-1141 multi) on the fused kernels.

Kept from the reference: alpha ramp (0.5 -> cap), beta = 1/mean-diff of the previous epoch,
gamma = 1e-3, lr clip schedule, 6 steps of 10 000 pairs (per environment) per epoch, the
5-deep (network, optimiser) snapshot ring and the "loss ratio not in (0, 1.2) -> reload a random
snapshot and redo the epoch" rule.  Snapshots are device-to-device copies of the flat buffers.
Under torch.distributed every rank processes its slice of each step's pairs and the flat gradient
is all-reduced; the rollback decision uses the all-reduced loss and a shared random stream.

Data feeding: every environment group's tables are stacked once into one
GPU-resident [rows, 2d+2] table; a step's batch is an int32 index vector (`perm[lo:hi] + row offset` per
environment) handed to pntf_loss_bwd_gather, so the reference's per-step `data[perm[lo:hi]]` gather, `stack`
and `contiguous` copies (M:1053-1066) never materialise.
"""
import random

import numpy as np
import torch

from . import capi
from .dist import GradAllReduce, shard_range, world


def run_training(model, datas, Bs, epochs=None, log=print, inner_batch=10000, steps_per_epoch=6, seed=0):
    multi = model.variant == capi.LOSS_MULTI
    net, opt = model.network, model.optimizer
    fp = net.flat()
    dev = fp.device
    d = model.dim
    rank, ws = world()
    n_epochs = int(epochs if epochs is not None else model.Params['Training']['Number of Epochs'])
    rng = random.Random(seed)                    # identical on every rank
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    reducer = GradAllReduce(capi.PARAM_COUNT, dev)

    env_batch = int(model.Params['Training']['Batch Size']) if multi else 1
    groups = [list(range(i, min(i + env_batch, len(datas)))) for i in range(0, len(datas), env_batch)]
    if multi:
        n_loader = len(groups)
        norm = n_loader * 5.0                    # M:1093-1094 divides by len(dataloader)*5 with 6 steps
    else:
        n_loader = int(np.ceil(datas[0].shape[0] / inner_batch))
        norm = float(n_loader)                   # S:1083-1084

    # GPU-resident tables, one per environment group (rows of the group's environments stacked)
    tables = []
    for grp in groups:
        tabs = [datas[e].to(device=dev, dtype=torch.float32) for e in grp]
        offs = [0]
        for t in tabs:
            offs.append(offs[-1] + t.shape[0])
        tables.append((torch.cat(tabs).contiguous(), offs))

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
        opt.param_groups[0]['lr'] = float(np.clip(1e-3 * (1 - (epoch - 8000) / 1000.), a_min=5e-4, a_max=1e-3))
        prev_diff = current_diff
        while True:
            totals = torch.zeros(2, dtype=torch.float32, device=dev)
            order = list(range(len(groups)))
            rng.shuffle(order)
            for gi in order:
                grp = groups[gi]
                E = len(grp)
                Ns = min(datas[e].shape[0] for e in grp)
                perms = [torch.randperm(Ns, device=dev, generator=gen) for _ in grp]
                B = torch.stack([Bs[e].to(torch.float32) for e in grp]).contiguous()
                nsteps = min(int(Ns / inner_batch), steps_per_epoch) if multi else min(n_loader, steps_per_epoch)
                for ii in range(max(nsteps, 1)):
                    lo, hi = ii * inner_batch, min((ii + 1) * inner_batch, Ns)
                    b, e_ = shard_range(hi - lo, rank, ws)
                    table, offs = tables[gi]
                    n = e_ - b
                    index = torch.cat([perms[j][lo + b:lo + e_] + offs[j] for j in range(E)]).to(torch.int32)
                    fp.grad.zero_()
                    loss_out = model._loss_step_gather(table, index, B, E, n, alpha, beta, gamma)
                    _, loss_out = reducer(fp.grad, loss_out)
                    opt.step(grad_scale=1.0 / ws)
                    totals += loss_out
            totals /= norm
            total_train_loss, total_diff = totals[0].item(), totals[1].item()
            current_diff = total_diff
            ratio = current_diff / prev_diff
            if 0 < ratio < 1.2:
                break
            k = rng.randint(0, 4)
            k = min(k, len(queue) - 1)        # the reference indexes past the ring before 5 epochs exist
            fp.flat.copy_(queue[k][0])
            opt.restore(queue[k][1])
            log("RepeatEpoch = {} -- Loss = {:.4e} -- Alpha = {:.4e}".format(epoch, total_diff, alpha))
        model.total_train_loss.append(total_train_loss)
        beta = 1.0 / total_diff
        history.append((epoch, total_diff, alpha))
        if rank == 0 and epoch % model.Params['Training']['Print Every * Epoch'] == 0:
            log("Epoch = {} -- Loss = {:.4e} -- Alpha = {:.4e}".format(epoch, total_diff, alpha))
        if rank == 0 and model.Params['ModelPath'] and (
                epoch % model.Params['Training']['Save Every * Epoch'] == 0 or epoch == n_epochs or epoch == 1):
            if multi:
                model.B = Bs[groups[order[-1]][0]]
            try:
                model.plot(epoch, total_diff, alpha)
                model.save(epoch=epoch, val_loss=total_diff)
            except OSError as exc:   # unwritable ModelPath: keep training, report
                log("checkpoint not written: %s" % exc)
    return history
