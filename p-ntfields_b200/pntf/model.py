"""Model classes with the reference's host-side API (constructor, Loss, train, save/load,
TravelTimes/Tau/Speed/Speed2/Gradient) plus a batched ``plan``.

Model_single mirrors models/model_res_sigmoid.py:829-1329, Model_multi mirrors
models/model_res_sigmoid_multi.py:857-1293.  The arithmetic runs in libpntf_b200.so; what stays
here is the trainer's control flow (alpha/beta/gamma/lr schedules, 5-deep rollback ring), the
checkpoint format and argument plumbing.
"""
import os
import random

import numpy as np
import torch

from . import capi
from .dist import gather_rows, shard_rows, world
from .nn import NN_multi, NN_single, _LossFn, _PREC
from .params import TENSOR_NAMES


def _default_params(ModelPath, DataPath, device, batch_size):
    P = {}
    P['ModelPath'] = ModelPath
    P['DataPath'] = DataPath
    P['Device'] = device
    P['Pytorch Amp (bool)'] = False
    P['Network'] = {'Normlisation': 'OffsetMinMax'}
    P['Training'] = {
        'Number of sample points': 2e5, 'Batch Size': batch_size, 'Validation Percentage': 10,
        'Number of Epochs': 10000, 'Resampling Bounds': [0.1, 0.9], 'Print Every * Epoch': 1,
        'Save Every * Epoch': 100, 'Learning Rate': 1e-3, 'Random Distance Sampling': True,
        'Use Scheduler (bool)': False}
    return P


class _FusedAdamW:
    """AdamW(lr=1e-3, weight_decay=0.1) on the flat parameter buffer with one fused kernel
    (reference: torch.optim.AdamW, S:954-956 / M:960-962).  Exposes param_groups[0]['lr'] and a
    torch-shaped state_dict so checkpoints keep the reference's keys.

    It reads the flat gradient buffer `fp.grad`, which the parameters' `.grad` alias (nn._NNBase.attach_grads), so
    whatever was done to `p.grad` between backward() and step() (scaling, clipping) is what gets applied.  One
    difference from torch.optim.AdamW remains: step() always updates every used parameter -- with an all-zero
    gradient buffer (no backward() since zero_grad()) that is a pure weight-decay / moment-decay step, where torch
    skips parameters whose .grad is None."""

    def __init__(self, network, lr=1e-3, weight_decay=0.1, betas=(0.9, 0.999), eps=1e-8):
        self.network = network
        self.param_groups = [{'lr': lr, 'weight_decay': weight_decay, 'betas': betas, 'eps': eps}]
        fp = network.flat(require_cuda=False)
        self.exp_avg = torch.zeros_like(fp.flat)
        self.exp_avg_sq = torch.zeros_like(fp.flat)
        self.step_count = 0

    def zero_grad(self, set_to_none=False):
        """Zeroes the flat gradient buffer (the parameters' .grad are views of it and stay attached)."""
        self.network.flat().grad.zero_()

    def step(self, grad_scale=1.0):
        fp = self.network.flat()
        g = self.param_groups[0]
        self.step_count += 1
        capi.adamw(fp.flat, fp.grad, self.exp_avg, self.exp_avg_sq, fp.mask, self.step_count, lr=g['lr'],
                   betas=g['betas'], eps=g['eps'], weight_decay=g['weight_decay'], grad_scale=grad_scale)

    def step_sched(self, lr_step, grad_scale=1.0):
        """step() with {lr, step count} taken from the device tensor lr_step (CUDA-graph replay; the caller advances
        lr_step[1] on the device and keeps step_count in sync after the replay)."""
        fp = self.network.flat()
        g = self.param_groups[0]
        capi.adamw_sched(fp.flat, fp.grad, self.exp_avg, self.exp_avg_sq, fp.mask, lr_step, betas=g['betas'],
                         eps=g['eps'], weight_decay=g['weight_decay'], grad_scale=grad_scale)

    def state_dict(self):
        fp = self.network.flat(require_cuda=False)
        state = {}
        idx = 0
        for k in TENSOR_NAMES:
            o, shp = fp.offsets[k], fp.shapes[k]
            n = int(np.prod(shp))
            if fp.mask[o].item() != 0 and self.step_count > 0:
                state[idx] = {'step': torch.tensor(float(self.step_count)),
                              'exp_avg': self.exp_avg[o:o + n].view(shp).clone(),
                              'exp_avg_sq': self.exp_avg_sq[o:o + n].view(shp).clone()}
            idx += 1
        g = dict(self.param_groups[0])
        g['params'] = list(range(len(TENSOR_NAMES)))
        return {'state': state, 'param_groups': [g]}

    def snapshot(self):
        return (self.exp_avg.clone(), self.exp_avg_sq.clone(), self.step_count)

    def restore(self, snap):
        self.exp_avg.copy_(snap[0])
        self.exp_avg_sq.copy_(snap[1])
        self.step_count = snap[2]


class _ModelBase:
    variant = None

    def _dev(self):
        return torch.device(self.Params['Device'])

    def gradient(self, y, x, create_graph=True):
        raise RuntimeError("autograd through the fused kernels w.r.t. coordinates is not available; use "
                           "network.out_grad / Model.Gradient (forward-mode derivatives computed on chip)")

    # -- checkpointing (S:1139-1162, M:1143-1166) --------------------------------------------------
    def save(self, epoch='', val_loss=''):
        torch.save({'epoch': epoch,
                    'model_state_dict': self.network.state_dict(),
                    'optimizer_state_dict': self.optimizer.state_dict(),
                    'B_state_dict': self.B,
                    'train_loss': self.total_train_loss,
                    'val_loss': self.total_val_loss},
                   '{}/Model_Epoch_{}_ValLoss_{:.6e}.pt'.format(self.Params['ModelPath'], str(epoch).zfill(5),
                                                                float(val_loss)))

    def load_pretrained_state_dict(self, state_dict):
        own_state = self.state_dict  # noqa: F841  (kept: the reference defines this stub, S:1165-1166)

    def _loss_eval(self, pts, yo, B, E, N, beta, gamma):
        """Loss() under torch.no_grad(): forward-only kernels (no stash, no parameter gradients)."""
        net = self.network
        loss_out, diff = capi.loss(self.variant, net.flat().flat, B, pts, yo, E, N, self.dim, 1.0, float(beta),
                                   float(gamma), prec=_PREC[net.precision])
        return loss_out[0], loss_out[1], diff

    # -- shared pieces of the trainer ------------------------------------------------------------------
    def _loss_step(self, coords, speed, B, E, N, alpha, beta, gamma):
        """Fused loss + gradient into the flat grad buffer; returns the device tensor [loss, loss_n]."""
        fp = self.network.flat()
        loss_out, _ = capi.loss(self.variant, fp.flat, B, coords, speed, E, N, self.dim, alpha, beta, gamma,
                                grad_out=fp.grad, prec=_PREC[self.network.precision],
                                budget=self.network.stash_budget, loss_out=fp.loss_slot)
        return loss_out

    def _loss_step_gather(self, table, index, B, E, N, alpha, beta, gamma):
        """_loss_step on the batch table[index] (GPU-resident dataset, gather fused into the kernel loads)."""
        fp = self.network.flat()
        loss_out, _ = capi.loss_gather(self.variant, fp.flat, B, table, index, E, N, self.dim, alpha, beta, gamma,
                                       grad_out=fp.grad, prec=_PREC[self.network.precision],
                                       budget=self.network.stash_budget, loss_out=fp.loss_slot,
                                       diff=self._diff_buffer(E * N))
        return loss_out

    def _diff_buffer(self, n):
        buf = getattr(self, "_diff_buf", None)
        if buf is None or buf.numel() < n or buf.device != self._dev():
            buf = torch.empty(n, dtype=torch.float32, device=self._dev())
            self._diff_buf = buf
        return buf[:n]

    # -- query sharding: planning and inference split their rows over the ranks of an initialised
    # torch.distributed group, one contiguous slice per rank (dist.shard_range), no collective.  shard=True returns this
    # rank's rows only; gather=True additionally all-gathers them so that every rank holds the whole result.
    def _rows(self, t, shard):
        if shard and world()[1] > 1:
            return shard_rows(t)[0]
        return t

    def _out(self, local, n_total, shard, gather):
        if shard and gather and world()[1] > 1:
            return gather_rows(local, n_total)
        return local

    def plan(self, start_goal, B=None, step=0.03, thr=0.06, max_iter=500, grad_mode=None, fill_tail=True, shard=False,
             gather=False):
        """Batched bidirectional path integration (test/gib_plan.py:53-97, test/arm_plan.py:111-158).

        start_goal[Q,2d] -> (waypoints[Q,max_iter+2,2d], nsteps[Q]); waypoints[q,:nsteps[q]+1,:d] is the
        start-side trajectory, waypoints[q,:nsteps[q]+1,d:] the goal-side one (``path()`` joins them).
        grad_mode: 'exact' or 'compat' (default: what the reference model uses -- see Gradient).
        fill_tail=False: rows after nsteps[q] are left unwritten (30x fewer bytes at the reference's cap of 500).
        shard / gather: split the queries over the ranks of torch.distributed (see above)."""
        if grad_mode is None:
            grad_mode = self.default_grad_mode
        if B is None:
            B = self.B if self.variant == capi.LOSS_MULTI else self.network.B
        dev = self._dev()
        Q = start_goal.shape[0]
        sg = self._rows(start_goal, shard).to(device=dev, dtype=torch.float32).contiguous()
        Bd = B.to(device=dev, dtype=torch.float32).contiguous()
        flav = capi.GRAD_COMPAT if grad_mode == 'compat' else capi.GRAD_EXACT
        way, nst = capi.plan(self.network.flat().flat, Bd, sg, self.dim, step, thr, max_iter, flav,
                             prec=_PREC[self.network.precision], fill_tail=fill_tail)
        return self._out(way, Q, shard, gather), self._out(nst, Q, shard, gather)

    @staticmethod
    def path(waypoints, nsteps, q, dim):
        """Start-side waypoints followed by the reversed goal-side waypoints (gib_plan.py:92-97)."""
        n = int(nsteps[q]) + 1
        w = waypoints[q, :n]
        return torch.cat([w[:, :dim], torch.flip(w[:, dim:], dims=[0])], dim=0)

    def _slice_points(self, Xsrc, limit, n):
        """(X, Y, XP): the multi model's slice (M:1250-1266): source (-0.25, -0.25, 0, ...), goal = the grid point in
        the first two coordinates and 0 beyond them.  Model_single overrides it with the UR5 slice of S:1284-1303."""
        spacing = 2 * limit / n
        ax = np.arange(-limit, limit, spacing)
        X, Y = np.meshgrid(ax, ax)
        src = [0.0] * self.dim if Xsrc is None else list(Xsrc)
        if Xsrc is None:
            src[0] = -0.25
            src[1] = -0.25
        XP = np.zeros((X.size, 2 * self.dim))
        XP[:, :self.dim] = src
        XP[:, self.dim + 0] = X.flatten()
        XP[:, self.dim + 1] = Y.flatten()
        return X, Y, XP

    def field_slice(self, Xsrc=None, limit=None, n=80):
        """Headless version of Model.plot (S:1284-1329, M:1250-1293): evaluates TravelTimes, Speed and
        Tau on an n x n slice and returns (X, Y, TT, V, TAU) as numpy arrays (no matplotlib)."""
        X, Y, XP = self._slice_points(Xsrc, self.slice_limit if limit is None else limit, n)
        XP = torch.tensor(XP, dtype=torch.float32, device=self._dev())
        tt = self.TravelTimes(XP)
        ss = self.Speed(XP)
        tau = self.Tau(XP)
        return (X, Y, tt.cpu().numpy().reshape(X.shape), ss.cpu().numpy().reshape(X.shape),
                tau.cpu().numpy().reshape(X.shape))

    def plot(self, epoch, total_train_loss, alpha):
        """The reference draws two JPEGs with matplotlib; headless here: arrays are saved as .npz."""
        X, Y, TT, V, TAU = self.field_slice()
        np.savez(os.path.join(self.Params['ModelPath'], "plots%s_%s_%s_0.npz" % (epoch, alpha, round(total_train_loss, 4))),
                 X=X, Y=Y, TT=TT, V=V, TAU=TAU)


class Model_single(_ModelBase):
    """Model(ModelPath, DataPath, dim, device) of models/model_res_sigmoid.py:829-1329 (UR5, d=6)."""
    variant = capi.LOSS_SINGLE
    default_grad_mode = 'exact'      # S:1254-1255 uses autograd: the exact gradient
    slice_limit = 2.0                # S:1286

    def _slice_points(self, Xsrc, limit, n):
        """S:1284-1303: a UR5 configuration (radians) scaled by 0.5 / pi as start AND goal, the goal's first two joints
        swept over [-limit, limit)^2."""
        spacing = 2 * limit / n
        ax = np.arange(-limit, limit, spacing)
        X, Y = np.meshgrid(ax, ax)
        if Xsrc is None:
            Xsrc = [-1.3, 0.4 - 0.5 * np.pi, 1.1, 0.5 - 0.5 * np.pi, -0.5, 0.7][:self.dim] + [0.0] * max(0, self.dim - 6)
        scale = np.pi / 0.5
        XP = np.zeros((X.size, 2 * self.dim))
        XP[:, :self.dim] = Xsrc
        XP[:, self.dim:] = Xsrc
        XP = XP / scale
        XP[:, self.dim + 0] = X.flatten()
        XP[:, self.dim + 1] = Y.flatten()
        return X, Y, XP

    def __init__(self, ModelPath, DataPath, dim, device='cpu'):
        self.Params = _default_params(ModelPath, DataPath, device, 10000)
        self.dim = dim
        self.total_train_loss = []
        self.total_val_loss = []
        self.alpha_cap = 1.05

    def Loss(self, points, Yobs, beta, gamma):
        """(loss, loss_n, diff[N]) -- S:869-936.  `loss` carries autograd edges to the parameters."""
        net = self.network
        dev = self._dev()
        pts = points.detach().to(device=dev, dtype=torch.float32).contiguous()
        yo = Yobs.detach().to(device=dev, dtype=torch.float32).contiguous()
        B = net.B.to(dtype=torch.float32).contiguous()
        if not torch.is_grad_enabled():
            return self._loss_eval(pts, yo, B.view(1, self.dim, -1), 1, pts.shape[0], beta, gamma)
        params = [dict(net.named_parameters())[k] for k in TENSOR_NAMES]
        return _LossFn.apply(net, self.variant, pts, yo, B, 1, pts.shape[0], 1.0, float(beta), float(gamma), *params)

    def init_network(self, B=None):
        if B is None:
            B = 0.5 * torch.normal(0, 1, size=(128, self.dim))
        self.B = B
        self.network = NN_single(self.Params['Device'], self.dim, self.B)
        self.network.apply(self.network.init_weights)
        self.network.to(self.Params['Device'])
        self.optimizer = _FusedAdamW(self.network, lr=self.Params['Training']['Learning Rate'], weight_decay=0.1)

    def load(self, filepath):
        checkpoint = torch.load(filepath, map_location=torch.device(self.Params['Device']), weights_only=False)
        self.B = checkpoint['B_state_dict']
        self.network = NN_single(self.Params['Device'], self.dim, self.B)
        self.network.load_state_dict(checkpoint['model_state_dict'], strict=True)
        self.network.to(torch.device(self.Params['Device']))
        self.network.float()
        self.network.eval()

    def train(self, data=None, epochs=None, log=print):
        """Training loop of S:938-1137 on fused kernels.  `data` = tensor [Ns, 2d+2] (points | speed);
        default: models/data_mlp.py layout, DataPath/sampled_points.npy + speed.npy."""
        from .data import load_single
        from .trainer import run_training
        self.init_network()
        if data is None:
            data = load_single(self.Params['DataPath'])
        return run_training(self, [data.to(self._dev())], [self.network.B], epochs, log)

    # -- inference ---------------------------------------------------------------------------------
    def _field(self, kind, Xp, gamma=0.0, flavour=capi.GRAD_EXACT, shard=False, gather=False):
        n = Xp.shape[0]
        Xp = self._rows(Xp.detach(), shard).to(device=self._dev(), dtype=torch.float32).contiguous()
        B = self.network.B.to(dtype=torch.float32).contiguous()
        out = capi.field(kind, self.network.flat().flat, B, Xp, 1, Xp.shape[0], self.dim, gamma, grad_flavour=flavour,
                         prec=_PREC[self.network.precision])
        return self._out(out, n, shard, gather)

    def TravelTimes(self, Xp, shard=False, gather=False):
        return self._field(capi.FIELD_TRAVEL_TIME, Xp, shard=shard, gather=gather)

    def Tau(self, Xp, shard=False, gather=False):
        n = Xp.shape[0]
        Xp = self._rows(Xp, shard).to(self._dev())
        tau, _ = self.network.out(Xp)
        return self._out(tau, n, shard, gather)

    def Speed(self, Xp, shard=False, gather=False):
        return self._field(capi.FIELD_SPEED, Xp, shard=shard, gather=gather)

    def Speed2(self, Xp, gamma, shard=False, gather=False):
        return self._field(capi.FIELD_SPEED2, Xp, gamma, shard=shard, gather=gather)

    def Gradient(self, Xp, shard=False, gather=False):
        """S:1247-1282.  The reference is only correct for batch 1; this is its row-wise extension
        (identical for one row)."""
        return self._field(capi.FIELD_DIRECTION, Xp, shard=shard, gather=gather)


class Model_multi(_ModelBase):
    """Model(ModelPath, DataPath, dim, length, device) of models/model_res_sigmoid_multi.py:857-1293."""
    variant = capi.LOSS_MULTI
    default_grad_mode = 'compat'     # M:1221 uses out_backgrad as shipped
    slice_limit = 0.5                # M:1251

    def __init__(self, ModelPath, DataPath, dim, length, device='cpu'):
        self.Params = _default_params(ModelPath, DataPath, device, 2)
        self.dim = dim
        self.len = length
        self.total_train_loss = []
        self.total_val_loss = []
        self.alpha_cap = 1.07
        self.B = None

    def Loss(self, points, Yobs, B, beta, gamma):
        """(loss, loss_n, diff[E,N]) -- M:897-951; points[E,N,2d], Yobs[E,N,2], B[E,d,128]."""
        net = self.network
        dev = self._dev()
        E, N = points.shape[0], points.shape[1]
        pts = points.detach().to(device=dev, dtype=torch.float32).contiguous().view(E * N, 2 * self.dim)
        yo = Yobs.detach().to(device=dev, dtype=torch.float32).contiguous().view(E * N, 2)
        Bd = B.detach().to(device=dev, dtype=torch.float32).contiguous()
        if not torch.is_grad_enabled():
            loss, loss_n, diff = self._loss_eval(pts, yo, Bd, E, N, beta, gamma)
            return loss, loss_n, diff.view(E, N)
        params = [dict(net.named_parameters())[k] for k in TENSOR_NAMES]
        loss, loss_n, diff = _LossFn.apply(net, self.variant, pts, yo, Bd, E, N, 1.0, float(beta), float(gamma), *params)
        return loss, loss_n, diff.view(E, N)

    def init_network(self):
        self.network = NN_multi(self.Params['Device'], self.dim)
        self.network.apply(self.network.init_weights)
        self.network.to(self.Params['Device'])
        self.optimizer = _FusedAdamW(self.network, lr=self.Params['Training']['Learning Rate'], weight_decay=0.1)

    def load(self, filepath):
        checkpoint = torch.load(filepath, map_location=torch.device(self.Params['Device']), weights_only=False)
        self.network = NN_multi(self.Params['Device'], self.dim)
        self.network.load_state_dict(checkpoint['model_state_dict'], strict=True)
        self.network.to(torch.device(self.Params['Device']))
        self.network.float()
        self.network.eval()

    def train(self, envs=None, epochs=None, log=print):
        """Training loop of M:953-1141 on fused kernels.  `envs` = list of (data[Ns,2d+2], B[d,128]);
        default: models/data_multi.py layout under DataPath."""
        from .data import load_multi
        from .trainer import run_training
        self.init_network()
        if envs is None:
            envs = load_multi(self.Params['DataPath'], self.len)
        datas = [e[0].to(self._dev()) for e in envs]
        Bs = [e[1].to(self._dev()) for e in envs]
        return run_training(self, datas, Bs, epochs, log)

    # -- inference (reads self.B like the reference, M:1173-1216) --------------------------------------
    def _field(self, kind, Xp, B, gamma=0.0, flavour=capi.GRAD_EXACT, shard=False, gather=False):
        n = Xp.shape[0]
        Xp = self._rows(Xp.detach(), shard).to(device=self._dev(), dtype=torch.float32).contiguous()
        Bd = B.detach().to(device=self._dev(), dtype=torch.float32).contiguous()
        out = capi.field(kind, self.network.flat().flat, Bd, Xp, 1, Xp.shape[0], self.dim, gamma, grad_flavour=flavour,
                         prec=_PREC[self.network.precision])
        return self._out(out, n, shard, gather)

    def TravelTimes(self, Xp, shard=False, gather=False):
        return self._field(capi.FIELD_TRAVEL_TIME, Xp, self.B, shard=shard, gather=gather)

    def Tau(self, Xp, shard=False, gather=False):
        n = Xp.shape[0]
        Xp = self._rows(Xp, shard).to(self._dev())
        tau, _ = self.network.out(Xp, self.B)
        return self._out(tau, n, shard, gather)

    def Speed(self, Xp, shard=False, gather=False):
        return self._field(capi.FIELD_SPEED, Xp, self.B, shard=shard, gather=gather)

    def Gradient(self, Xp, B, grad_mode='compat', shard=False, gather=False):
        """M:1218-1248.  grad_mode='compat' (default) reproduces the reference as shipped (its
        out_backgrad evaluates the first encoder gate at softplus(y), M:435-438); 'exact' uses the
        true gradient of tau."""
        flav = capi.GRAD_COMPAT if grad_mode == 'compat' else capi.GRAD_EXACT
        return self._field(capi.FIELD_DIRECTION, Xp, B, flavour=flav, shard=shard, gather=gather)
