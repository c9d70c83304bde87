"""NN classes with the reference's method names and signatures, computed by libpntf_b200.so.

NN_single mirrors models/model_res_sigmoid.py:129-826  (NN(device, dim, B); B baked in)
NN_multi  mirrors models/model_res_sigmoid_multi.py:129-854 (NN(device, dim); B passed per call)

Both are torch.nn.Module objects holding real ``torch.nn.Linear`` sub-modules under the reference's
names (encoder / encoder1 / generator / generator1), so ``state_dict()`` keys and shapes, ``apply``,
``parameters()``, ``load_state_dict(strict=True)`` and ``torch.save`` behave exactly like the
reference.  The parameter tensors are views into ONE flat fp32 buffer (FlatParams) which is what the
kernels read; ``flat()`` re-packs if a caller replaced parameter storage (e.g. ``.to(device)``).
"""
import math

import torch
from torch.nn import Linear

from . import capi
from .params import FlatParams, TENSOR_NAMES, UNUSED

_PREC = {"fp32": capi.PREC_FP32, "tc": capi.PREC_TC}


class _LossFn(torch.autograd.Function):
    """loss = beta * loss_n with gradients w.r.t. the parameters.

    forward runs pntf_loss_bwd (loss AND d loss / d params in one pass) into a scratch buffer; backward adds
    ``g * scratch`` to the flat gradient buffer ``fp.grad`` (g = the incoming gradient of `loss`, plus that of
    `loss_n` divided by beta), which is the buffer the fused optimiser reads and which every parameter's ``.grad``
    aliases (views).  So ``(2 * loss).backward()``, several backward calls before one ``step()``, gradient clipping on
    ``p.grad`` and ``zero_grad()`` behave as with the reference's torch.optim.AdamW; no per-step allocation.
    The parameter inputs receive no autograd gradient of their own (``torch.autograd.grad(loss, params)`` is not
    supported: read ``p.grad`` after ``backward()``)."""

    @staticmethod
    def forward(ctx, net, variant, coords, speed, B, E, N, alpha, beta, gamma, *params):
        fp = net.flat()
        tmp = fp.scratch_grad()
        tmp.zero_()
        loss_out, diff = capi.loss(variant, fp.flat, B, coords, speed, E, N, net.dim, alpha, beta, gamma,
                                   grad_out=tmp, prec=_PREC[net.precision], budget=net.stash_budget)
        ctx.net = net
        ctx.beta = float(beta)
        ctx.mark_non_differentiable(diff)
        return loss_out[0], loss_out[1], diff

    @staticmethod
    def backward(ctx, g_loss, g_loss_n, g_diff):
        net = ctx.net
        fp = net.flat()
        g = g_loss
        if g_loss_n is not None and ctx.beta != 0.0:
            g = g_loss_n / ctx.beta if g is None else g + g_loss_n / ctx.beta
        if g is not None:
            tmp = fp.scratch_grad()
            fp.grad.addcmul_(tmp, g.to(tmp.dtype).expand_as(tmp))
        net.attach_grads()
        return (None,) * (10 + len(TENSOR_NAMES))


class _NNBase(torch.nn.Module):
    def __init__(self, device, dim):
        super().__init__()
        self.dim = dim
        self.scale = 10
        self.nl1 = 3
        self.nl2 = 3
        h = 128
        self.input_size = 128
        self.encoder = torch.nn.ModuleList()
        self.encoder1 = torch.nn.ModuleList()
        self.encoder.append(Linear(2 * h, h))
        self.encoder1.append(Linear(2 * h, h))
        for _ in range(self.nl1 - 1):
            self.encoder.append(Linear(h, h))
            self.encoder1.append(Linear(h, h))
        self.encoder.append(Linear(h, h))
        self.generator = torch.nn.ModuleList()
        self.generator1 = torch.nn.ModuleList()
        for _ in range(self.nl2):
            self.generator.append(Linear(2 * h, 2 * h))
            self.generator1.append(Linear(2 * h, 2 * h))
        self.generator.append(Linear(2 * h, h))
        self.generator.append(Linear(h, 1))
        self._fp = None
        self._device = torch.device(device)
        #: 'tc' = tcgen05 tensor-core mode (fp16 operands, fp32 accumulation), 'fp32' = fp32 SIMT verify mode
        self.precision = "tc"
        #: bytes of activation stash pntf_loss_bwd may use (0 = library default)
        self.stash_budget = 0

    # -- reference API -------------------------------------------------------------------------
    def init_weights(self, m):
        """U(+-2/sqrt(fan_in)) on weight and bias of every Linear (S:183-189, M:177-182)."""
        if type(m) == torch.nn.Linear:
            stdv = (1. / math.sqrt(m.weight.size(1)) / 1.) * 2
            m.weight.data.uniform_(-stdv, stdv)
            m.bias.data.uniform_(-stdv, stdv)

    # -- flat parameter plumbing -----------------------------------------------------------------
    def flat(self, require_cuda=True) -> FlatParams:
        """The flat parameter buffer the kernels read; parameters are views into it.  require_cuda=False is for
        bookkeeping only (state_dict / checkpoints of a network that lives on the CPU); every compute path asks for
        a CUDA buffer and raises otherwise: there is no CPU fallback."""
        named = dict(self.named_parameters())
        dev = next(iter(named.values())).device
        if require_cuda and dev.type != "cuda":
            raise RuntimeError("pntf NN needs its parameters on a CUDA device (no CPU fallback); got %s" % dev)
        fp = self._fp
        ok = fp is not None and fp.device == dev
        if ok:
            base = fp.flat.data_ptr()
            for k in TENSOR_NAMES:
                if named[k].data_ptr() != base + 4 * fp.offsets[k]:
                    ok = False
                    break
        if not ok:
            fp = FlatParams(dev)
            with torch.no_grad():
                for k in TENSOR_NAMES:
                    fp.views[k].copy_(named[k].data.to(torch.float32))
                    named[k].data = fp.views[k]
            self._fp = fp
        return fp

    def attach_grads(self):
        """Every used parameter's .grad is a view of the flat gradient buffer (encoder1.0.* stays None: the reference
        never gives it a gradient, S:166 vs S:225)."""
        fp = self.flat()
        for k, p in self.named_parameters():
            if k in UNUSED:
                continue
            v = fp.grad_views[k]
            if p.grad is None or p.grad.data_ptr() != v.data_ptr():
                p.grad = v

    def _run(self, mode, coords, B, E, N, flavour=capi.GRAD_EXACT):
        coords = coords.detach().to(dtype=torch.float32).contiguous()
        B = B.detach().to(device=coords.device, dtype=torch.float32).contiguous()
        fp = self.flat()
        return capi.forward(mode, fp.flat, B, coords.view(E * N, 2 * self.dim), E, N, self.dim,
                            prec=_PREC[self.precision],
                            grad_flavour=flavour)


class NN_single(_NNBase):
    """Single-environment network: NN(device, dim, B) with B[128, dim] (S:131-181)."""

    def __init__(self, device, dim, B):
        super().__init__(device, dim)
        self.B = B.T.to(device)          # [dim, 128] as the reference stores it (S:139)

    def out(self, coords):
        tau, _, _ = self._run(capi.MODE_VALUE, coords, self.B, 1, coords.shape[0])
        return tau.view(-1, 1), coords

    def out_grad(self, coords):
        tau, dtau, _ = self._run(capi.MODE_GRAD, coords, self.B, 1, coords.shape[0])
        return tau.view(-1, 1), dtau, coords

    def out_laplace(self, coords):
        tau, dtau, ltau = self._run(capi.MODE_LAPLACE, coords, self.B, 1, coords.shape[0])
        return tau.view(-1, 1), dtau, ltau, coords

    def forward(self, coords):
        return self.out(coords)


class NN_multi(_NNBase):
    """Multi-environment network: NN(device, dim); B[d,128] (or [E,d,128]) passed per call (M:131-175)."""

    def __init__(self, device, dim):
        super().__init__(device, dim)

    def out(self, coords, B):
        tau, _, _ = self._run(capi.MODE_VALUE, coords, B, 1, coords.shape[0])
        return tau.view(-1, 1), coords

    def out_grad(self, coords, B):
        tau, dtau, _ = self._run(capi.MODE_GRAD, coords, B, 1, coords.shape[0])
        return tau.view(-1, 1), dtau, coords

    def out_backgrad(self, coords, B):
        """The reference's hand-written reverse sweep (M:402-647) INCLUDING its first-layer gate
        evaluated at softplus(y) (M:435-438): this is the 'compat' gradient flavour."""
        tau, dtau, _ = self._run(capi.MODE_GRAD, coords, B, 1, coords.shape[0], capi.GRAD_COMPAT)
        return tau.view(-1, 1), dtau, coords

    def out_laplace(self, coords, B):
        E, N = coords.shape[0], coords.shape[1]
        tau, dtau, ltau = self._run(capi.MODE_LAPLACE, coords, B, E, N)
        d2 = 2 * self.dim
        return tau.view(E, N, 1), dtau.view(E, N, d2), ltau.view(E, N, d2), coords

    def forward(self, coords, B):
        return self.out(coords, B)
