"""CPU oracle for the P-NTFields hot path.  TEST INFRASTRUCTURE ONLY.

This file is a from-scratch CPU restatement (PyTorch CPU tensors, fp32 or fp64)
of the arithmetic on the reference's hot path.  It is the *checker*: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it.  Nothing under ``p-ntfields_b200/``
imports it, and the product path fails loudly when the CUDA library is missing.

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the unmodified
reference (``/root/reference/models/model_res_sigmoid{,_multi}.py``) in the
build container, runs it on seeded inputs and commits its outputs as
``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every function
below against those fixtures (the reference itself ships no tests or golden
vectors).

Notation (S: = models/model_res_sigmoid.py, M: = models/model_res_sigmoid_multi.py):
  d        C-space dimension, a start/goal pair is a row (x0[d], x1[d])
  column   one of: value, d first-derivative columns, d second-derivative columns
           (per endpoint in the encoder; 2d + 2d per pair in the generator)
All tensors that carry columns are laid out [rows, columns, features].
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

H = 128          # hidden width of the encoder (S:135, M:133)
NFREQ = 128      # Fourier frequencies (S:141, M:134)
SCALE = 10.0     # softplus / sigmoid sharpness (S:144, M:137)

# (name, out_features, in_features) in state_dict order (S:159-181, M:153-175).
LAYER_TABLE = (
    ("encoder.0", 128, 256), ("encoder.1", 128, 128), ("encoder.2", 128, 128),
    ("encoder.3", 128, 128),
    ("encoder1.0", 128, 256),   # allocated, never used by any forward (S:166 vs S:225)
    ("encoder1.1", 128, 128), ("encoder1.2", 128, 128),
    ("generator.0", 256, 256), ("generator.1", 256, 256), ("generator.2", 256, 256),
    ("generator.3", 128, 256), ("generator.4", 1, 128),
    ("generator1.0", 256, 256), ("generator1.1", 256, 256), ("generator1.2", 256, 256),
)


def init_params(seed: int = 0, dtype=torch.float32) -> "OrderedDict[str, torch.Tensor]":
    """Random-init weights: W and b ~ U(+-2/sqrt(fan_in)) (S:183-189, M:177-182).

    Drawn from numpy's frozen legacy RandomState (layer by layer, weight then bias) so the
    fixtures under tests/golden stay valid across torch versions; it reproduces the
    distribution of the reference init, not torch's RNG consumption.
    """
    import numpy as np
    rs = np.random.RandomState(seed)
    sd = OrderedDict()
    for name, fo, fi in LAYER_TABLE:
        bound = 2.0 / math.sqrt(fi)
        sd[name + ".weight"] = torch.from_numpy(rs.uniform(-bound, bound, size=(fo, fi))).to(dtype)
        sd[name + ".bias"] = torch.from_numpy(rs.uniform(-bound, bound, size=(fo,))).to(dtype)
    return sd


# --------------------------------------------------------------------------------------
# activations (S:75-126, S:146; M:76-126, M:139)
# --------------------------------------------------------------------------------------
def softplus10(y: torch.Tensor) -> torch.Tensor:
    """a(y) = log(1+exp(10y))/10, torch's threshold rule: returns y when 10y > 20."""
    return torch.nn.functional.softplus(y, beta=SCALE, threshold=20.0)


def sig10(y: torch.Tensor) -> torch.Tensor:
    return torch.sigmoid(SCALE * y)


def _act_cols(t: torch.Tensor, order: int, first_dact_of_act: bool = False) -> torch.Tensor:
    """Hidden activation on a column stack t[rows, C, F] (S:641-657 / M:675-691 for order 2,
    S:279-287 for order 1).  Column 0 is the value, columns 1..n the first derivatives and
    (order 2) columns n+1..2n the matching diagonal second derivatives.

    first_dact_of_act reproduces the statement order of M:435-438 (out_backgrad's first
    encoder layer evaluates the derivative gate at softplus(y) instead of y).
    """
    y = t[:, :1, :]
    val = softplus10(y)
    s = sig10(val) if first_dact_of_act else sig10(y)
    if order == 0:
        return val
    ncol = t.shape[1] - 1
    n = ncol if order == 1 else ncol // 2
    dy = t[:, 1:1 + n, :]
    out = [val, dy * s]
    if order == 2:
        sp = SCALE * s * (1.0 - s)
        ly = t[:, 1 + n:, :]
        out.append(dy * dy * sp + ly * s)
    return torch.cat(out, dim=1)


def _actout_cols(t: torch.Tensor, order: int) -> torch.Tensor:
    """Output activation o(y)=sigmoid(0.1 y) with its derivative rules (S:659-674, S:289-297)."""
    y = t[:, :1, :]
    o = torch.sigmoid(0.1 * y)
    if order == 0:
        return o
    ncol = t.shape[1] - 1
    n = ncol if order == 1 else ncol // 2
    o1 = 0.1 * o * (1.0 - o)
    dy = t[:, 1:1 + n, :]
    out = [o, dy * o1]
    if order == 2:
        o2 = 0.1 * o1 * (1.0 - 2.0 * o)
        ly = t[:, 1 + n:, :]
        out.append(dy * dy * o2 + ly * o1)
    return torch.cat(out, dim=1)


def _linear_cols(t: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """One GEMM over all columns with a shared W; bias on the value column only (S:629-639)."""
    y = t @ w.T
    y[:, 0, :] = y[:, 0, :] + b
    return y


def fourier_cols(x: torch.Tensor, Bw: torch.Tensor, order: int) -> torch.Tensor:
    """Fourier features of endpoints x[R,d] with B[d,128] (S:191-210, M:186-213).

    Returns [R, 1+order*d, 256]: value [sin p, cos p], d/dx_i = [w_i cos p, -w_i sin p],
    d2/dx_i2 = [-w_i^2 sin p, -w_i^2 cos p], with w = 2*pi*B and p = x @ w.
    """
    w = (2.0 * math.pi) * Bw                      # [d, 128]
    p = x @ w                                     # [R, 128]
    sp, cp = torch.sin(p), torch.cos(p)
    cols = [torch.cat([sp, cp], dim=-1).unsqueeze(1)]
    if order >= 1:
        wi = w.unsqueeze(0)                       # [1, d, 128]
        cols.append(torch.cat([wi * cp.unsqueeze(1), -wi * sp.unsqueeze(1)], dim=-1))
    if order >= 2:
        cols.append(torch.cat([-wi * wi * sp.unsqueeze(1), -wi * wi * cp.unsqueeze(1)], dim=-1))
    return torch.cat(cols, dim=1)


# --------------------------------------------------------------------------------------
# network forward: tau, grad tau, diagonal second derivatives (forward mode)
# --------------------------------------------------------------------------------------
def _encoder(P: Dict[str, torch.Tensor], t: torch.Tensor, order: int, compat_first_dact: bool) -> torch.Tensor:
    """Per-endpoint encoder: L0, act, 2 x {L, act, L, +skip, act}, L (S:690-722, M:724-750)."""
    t = _act_cols(_linear_cols(t, P["encoder.0.weight"], P["encoder.0.bias"]), order, compat_first_dact)
    for ii in (1, 2):
        skip = t
        t = _act_cols(_linear_cols(t, P[f"encoder.{ii}.weight"], P[f"encoder.{ii}.bias"]), order)
        t = _linear_cols(t, P[f"encoder1.{ii}.weight"], P[f"encoder1.{ii}.bias"]) + skip
        t = _act_cols(t, order)
    return _linear_cols(t, P["encoder.3.weight"], P["encoder.3.bias"])


def _combine(f0: torch.Tensor, f1: torch.Tensor, order: int, d: int) -> torch.Tensor:
    """Symmetric smooth max / smooth min of the two endpoint feature stacks (S:724-787, M:752-813).

    f0, f1: [N, 1+order*d, 128].  Output [N, 1+order*2d, 256]; derivative columns are ordered
    start coordinates first, then goal coordinates.
    """
    v0, v1 = f0[:, 0, :], f1[:, 0, :]
    both = torch.stack([v0, v1], dim=1)
    smax = torch.logsumexp(SCALE * both, dim=1) / SCALE
    smin = -torch.logsumexp(-SCALE * both, dim=1) / SCALE
    cols = [torch.cat([smax, smin], dim=-1).unsqueeze(1)]
    if order >= 1:
        s0 = sig10(v0 - v1).unsqueeze(1)
        s1 = 1.0 - s0
        a, b = f0[:, 1:1 + d, :], f1[:, 1:1 + d, :]
        cols.append(torch.cat([torch.cat([a * s0, b * s1], dim=1),
                               torch.cat([a * s1, b * s0], dim=1)], dim=-1))
    if order >= 2:
        q = SCALE * s0 * s1
        A, Bq = f0[:, 1 + d:, :], f1[:, 1 + d:, :]
        qa, qb = q * a * a, q * b * b
        cols.append(torch.cat([torch.cat([qa + A * s0, qb + Bq * s1], dim=1),
                               torch.cat([-qa + A * s1, -qb + Bq * s0], dim=1)], dim=-1))
    return torch.cat(cols, dim=1)


def _generator(P: Dict[str, torch.Tensor], t: torch.Tensor, order: int) -> torch.Tensor:
    """3 x {L, act, L, +skip, act}, L(256->128), act, L(128->1), output act (S:790-814, M:816-840)."""
    for ii in (0, 1, 2):
        skip = t
        t = _act_cols(_linear_cols(t, P[f"generator.{ii}.weight"], P[f"generator.{ii}.bias"]), order)
        t = _linear_cols(t, P[f"generator1.{ii}.weight"], P[f"generator1.{ii}.bias"]) + skip
        t = _act_cols(t, order)
    t = _act_cols(_linear_cols(t, P["generator.3.weight"], P["generator.3.bias"]), order)
    return _actout_cols(_linear_cols(t, P["generator.4.weight"], P["generator.4.bias"]), order)


def forward(P: Dict[str, torch.Tensor], coords: torch.Tensor, Bw: torch.Tensor, order: int = 2,
            compat_first_dact: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor]]:
    """tau[N,1], dtau[N,2d] (order>=1), ltau[N,2d] (order 2) for pairs coords[N,2d], B[d,128].

    order 0 = NN.out (S:212-256), 1 = NN.out_grad (S:513-613), 2 = NN.out_laplace (S:676-820).
    compat_first_dact (order 1 only) reproduces the direction M:402-647 (out_backgrad) returns:
    its reverse sweep equals this forward-mode pass with the first encoder gate evaluated at
    softplus(y) (M:435-438); with False it is the exact gradient.
    """
    N, two_d = coords.shape
    d = two_d // 2
    x = torch.cat([coords[:, :d], coords[:, d:]], dim=0)          # [2N, d]: starts then goals
    t = fourier_cols(x, Bw, order)
    t = _encoder(P, t, order, compat_first_dact)
    t = _combine(t[:N], t[N:], order, d)
    t = _generator(P, t, order)[:, :, 0]                          # [N, 1+order*2d]
    tau = t[:, :1]
    dtau = t[:, 1:1 + 2 * d] if order >= 1 else None
    ltau = t[:, 1 + 2 * d:] if order >= 2 else None
    return tau, dtau, ltau


# --------------------------------------------------------------------------------------
# Eikonal residual loss (S:869-936 single; M:897-951 multi)
# --------------------------------------------------------------------------------------
def _slowness_terms(coords, tau, dtau, d):
    D = coords[:, d:] - coords[:, :d]
    T0 = (D * D).sum(-1)
    t = tau[:, 0]
    g0, g1 = dtau[:, :d], dtau[:, d:]
    S0 = T0 * (g0 * g0).sum(-1) + 2.0 * t * (g0 * D).sum(-1) + t * t
    S1 = T0 * (g1 * g1).sum(-1) - 2.0 * t * (g1 * D).sum(-1) + t * t
    return D, T0, t, S0, S1


def loss_single(P, coords, speed, Bw, beta: float, gamma: float):
    """Model.Loss of the single-environment model (S:869-936): returns (loss, loss_n, diff[N])."""
    d = coords.shape[1] // 2
    tau, dtau, ltau = forward(P, coords, Bw, 2)
    _, _, t, S0, S1 = _slowness_terms(coords, tau, dtau, d)
    diff = 0.0
    for k, S in enumerate((S0, S1)):
        lap = ltau[:, k * d:(k + 1) * d].sum(-1)
        visc = 1.0 / (torch.sqrt(S) / (t * t) + gamma * lap)      # 1/(1/Ypred + gamma*lap), S:913-918
        r = torch.sqrt(visc) / torch.sqrt(speed[:, k])
        diff = diff + r + 1.0 / r
    diff = diff - 4.0
    loss_n = diff.sum() / speed.shape[0]
    return beta * loss_n, loss_n, diff


def loss_multi(P, coords, speed, Bw, beta: float, gamma: float):
    """Model.Loss of the multi-environment model (M:897-951): coords[E,N,2d], speed[E,N,2],
    B[E,d,128]; no square roots on the ratio; + 0.01*|B|^2/(E*N) inside loss_n (M:947)."""
    E, N, two_d = coords.shape
    d = two_d // 2
    diffs = []
    for e in range(E):
        tau, dtau, ltau = forward(P, coords[e], Bw[e], 2)
        _, _, t, S0, S1 = _slowness_terms(coords[e], tau, dtau, d)
        df = 0.0
        for k, S in enumerate((S0, S1)):
            lap = ltau[:, k * d:(k + 1) * d].sum(-1)
            pred = 1.0 / (torch.sqrt(S) / (t * t) + gamma * lap)
            r = pred / speed[e, :, k]
            df = df + r + 1.0 / r
        diffs.append(df - 4.0)
    diff = torch.stack(diffs, dim=0)
    loss_n = diff.sum() / E / N + 0.01 * torch.linalg.norm(Bw) ** 2 / E / N
    return beta * loss_n, loss_n, diff


# --------------------------------------------------------------------------------------
# inference API (S:1169-1245, M:1173-1216)
# --------------------------------------------------------------------------------------
def travel_times(P, Xp, Bw):
    d = Xp.shape[1] // 2
    tau, _, _ = forward(P, Xp, Bw, 0)
    D = Xp[:, d:] - Xp[:, :d]
    return torch.sqrt((D * D).sum(-1)) / tau[:, 0]


def speed_pred(P, Xp, Bw, gamma: Optional[float] = None):
    """Model.Speed (goal side only, S:1192-1216); with gamma: Model.Speed2 (S:1218-1245)."""
    d = Xp.shape[1] // 2
    tau, dtau, ltau = forward(P, Xp, Bw, 1 if gamma is None else 2)
    _, _, t, _, S1 = _slowness_terms(Xp, tau, dtau, d)
    if gamma is None:
        return t * t / torch.sqrt(S1)
    return 1.0 / (torch.sqrt(S1) / (t * t) + gamma * ltau[:, d:].sum(-1))


def descent_direction(P, Xp, Bw, grad_mode: str = "exact"):
    """Model.Gradient, row-wise (M:1218-1248; S:1247-1282 is the batch-1 special case).

    grad_mode 'exact' uses the true grad tau (what S: gets from autograd, S:1254-1255);
    'compat' uses the direction M:402-647 actually returns (see forward()).
    """
    d = Xp.shape[1] // 2
    tau, dtau, _ = forward(P, Xp, Bw, 1, compat_first_dact=(grad_mode == "compat"))
    D = Xp[:, d:] - Xp[:, :d]
    T0 = torch.sqrt((D * D).sum(-1))
    t = tau[:, 0]
    c1 = (1.0 / (T0 * t)).unsqueeze(1)
    c2 = (T0 / (t * t)).unsqueeze(1)
    out = []
    for V0, V1 in ((-D, dtau[:, :d]), (D, dtau[:, d:])):
        Y = -(c1 * V0 - c2 * V1)
        nrm = torch.linalg.norm(Y, dim=1, keepdim=True)
        out.append(Y / nrm ** 2)
    return torch.cat(out, dim=1)


def plan(P, start_goal, Bw, step: float, thr: float, max_iter: int, grad_mode: str = "exact"):
    """Batched restatement of the plan loop (test/gib_plan.py:53-86, test/arm_plan.py:111-152):
    per query, while |x1-x0| > thr: XP += step*Gradient(XP); stop after iteration max_iter+1.

    Returns waypoints[Q, max_iter+2, 2d] (row 0 = the query; rows after a query's last step
    repeat its final position) and nsteps[Q].
    """
    Q, two_d = start_goal.shape
    d = two_d // 2
    XP = start_goal.clone()
    way = XP.unsqueeze(1).repeat(1, max_iter + 2, 1)
    nsteps = torch.zeros(Q, dtype=torch.int32)
    active = torch.linalg.norm(XP[:, d:] - XP[:, :d], dim=1) > thr
    it = 0
    while bool(active.any()):
        idx = active.nonzero()[:, 0]
        g = descent_direction(P, XP[idx], Bw, grad_mode)
        XP[idx] = XP[idx] + step * g
        it += 1
        nsteps[idx] = it
        way[idx, it:, :] = XP[idx].unsqueeze(1)
        dis = torch.linalg.norm(XP[:, d:] - XP[:, :d], dim=1)
        active = active & (dis > thr)
        if it > max_iter:
            break
    return way, nsteps


# --------------------------------------------------------------------------------------
# optimiser step (torch.optim.AdamW, lr 1e-3, weight_decay 0.1: S:954-956, M:960-962)
# --------------------------------------------------------------------------------------
def adamw_step(p, g, m, v, step: int, lr: float = 1e-3, b1: float = 0.9, b2: float = 0.999,
               eps: float = 1e-8, wd: float = 0.1):
    """One decoupled-weight-decay Adam update on flat tensors; returns (p, m, v)."""
    p = p * (1.0 - lr * wd)
    m = m + (g - m) * (1.0 - b1)
    v = v * b2 + (1.0 - b2) * g * g
    bc1 = 1.0 - b1 ** step
    bc2 = 1.0 - b2 ** step
    denom = torch.sqrt(v) / math.sqrt(bc2) + eps
    p = p - (lr / bc1) * m / denom
    return p, m, v


def loss_and_grads(P, coords, speed, Bw, beta, gamma, variant: str):
    """Loss + parameter gradients by autograd (the reference's loss.backward(), S:1071, M:1076).
    Returns (loss, loss_n, diff, grads{name: tensor or None})."""
    Pg = OrderedDict((k, v.detach().clone().requires_grad_(True)) for k, v in P.items())
    fn = loss_single if variant == "single" else loss_multi
    loss, loss_n, diff = fn(Pg, coords, speed, Bw, beta, gamma)
    loss.backward()
    grads = OrderedDict((k, v.grad) for k, v in Pg.items())
    return loss.detach(), loss_n.detach(), diff.detach(), grads
