"""Parity at the tolerances BASELINE.json's north_star states, POINTWISE, on the BASELINE workloads' own input
distribution: pairs drawn by the reference's sampler rule with starts kept inside the obstacle band
(dataprocessing/speed_sampling_gpu.py:346-365), speeds clipped at 0.1 next to obstacles (:415-419), coordinates rounded
through fp16 (models/data_multi.py:19).

  tensor-core mode : tau within rel 1e-3, speed and loss within rel 1e-2
  fp32 verify mode : tau, speed and loss within rel 1e-5
  TC backward      : parameter gradients after 200 optimiser steps against fp32 mode and oracle autograd, with the
                     fp16 ranges of the operand and adjoint tiles checked (no saturation, no flush)

Measured error quantiles (B200, this code; scripts/explore_parity.py) are quoted next to each bound so that a bound
wider than north_star's is visible as such.  Oracle = oracle/pntf_oracle.py in 10 000-pair chunks (the reference's
micro-batch), pinned to the reference by tests/test_oracle_golden.py.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
GAMMA = 1e-3


@pytest.fixture(scope="module")
def capi():
    from pntf import capi as c
    c.lib()
    torch.set_num_threads(os.cpu_count() or 1)
    return c


def _flat(P):
    return torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()


def _oracle(P, c, s, B, E, N, single, dtype=torch.float32, chunk=10000):
    """tau, dtau, speed (goal side), speed2, diff per pair, and loss_n, by the oracle in chunks."""
    from oracle import pntf_oracle as O
    Pd = {k: v.to(dtype) for k, v in P.items()}
    taus, dtaus, sps, sp2s, diffs = [], [], [], [], []
    with torch.no_grad():
        for e in range(E):
            Bw = B[e].to(dtype)
            for lo in range(0, N, chunk):
                cc = c[e * N + lo:e * N + min(lo + chunk, N)].to(dtype)
                ss = s[e * N + lo:e * N + min(lo + chunk, N)].to(dtype)
                tau, dtau, _ = O.forward(Pd, cc, Bw, 1)
                if single:
                    _, _, diff = O.loss_single(Pd, cc, ss, Bw, 1.0, GAMMA)
                else:
                    _, _, diff = O.loss_multi(Pd, cc.unsqueeze(0), ss.unsqueeze(0), Bw.unsqueeze(0), 1.0, GAMMA)
                    diff = diff[0]
                taus.append(tau[:, 0]); dtaus.append(dtau); diffs.append(diff)
                sps.append(O.speed_pred(Pd, cc, Bw))
                sp2s.append(O.speed_pred(Pd, cc, Bw, GAMMA))
    diff = torch.cat(diffs)
    loss_n = float(diff.double().sum() / (E * N))
    if not single:
        loss_n += 0.01 * float((B.double() ** 2).sum()) / (E * N)
    return torch.cat(taus), torch.cat(dtaus), torch.cat(sps), torch.cat(sp2s), diff, loss_n


def _rel(a, b, floor=0.0):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor)


def _workload(d, E, N, margin, seed=2):
    from pntf import synth
    B = synth.sample_B(d, DEV, seed=1, E=E).contiguous()
    c, s = synth.sample_pairs(E * N, d, DEV, seed=seed, margin=margin, near_obstacle=True)
    assert float(s.min()) <= 0.1 + 1e-6            # the clip floor is present in the batch
    return B, c, s


def _cuda_outputs(capi, flat, B, c, s, E, N, d, variant, prec):
    tau, dtau, _ = capi.forward(capi.MODE_LAPLACE, flat, B, c, E, N, d, prec=prec)
    lo, diff = capi.loss(variant, flat, B, c, s, E, N, d, 1.0, 1.0, GAMMA, prec=prec)
    sp, sp2 = [], []
    for e in range(E):
        Be, ce = B[e:e + 1].contiguous(), c[e * N:(e + 1) * N].contiguous()
        sp.append(capi.field(capi.FIELD_SPEED, flat, Be, ce, 1, N, d, prec=prec))
        sp2.append(capi.field(capi.FIELD_SPEED2, flat, Be, ce, 1, N, d, gamma=GAMMA, prec=prec))
    torch.cuda.synchronize()
    return tau.cpu(), dtau.cpu(), torch.cat(sp).cpu(), torch.cat(sp2).cpu(), diff.cpu(), float(lo[1])


@pytest.mark.parametrize("label,d,E,N,margin,single", [
    ("configs[0] test box, E=1 x 100 000", 3, 1, 100000, 0.5 / 5, False),
    ("configs[0] test box, E=2 x 50 000", 3, 2, 50000, 0.5 / 5, False),
    ("configs[3] UR5 d=6, 20 000-pair slice", 6, 1, 20000, 0.5 / 12, True)])
def test_tensor_core_mode_pointwise_on_the_baseline_workloads(capi, label, d, E, N, margin, single):
    """north_star: tau rel 1e-3, speed and loss rel 1e-2 (tensor-core mode).  Every bound below is POINTWISE.

    Measured maxima over the batch (d=3 100 000 pairs / d=6 20 000 pairs): tau 1.8e-4 / 1.6e-4; Speed 8.6e-4 / 7.6e-4;
    loss_n 7e-6 / 2e-5.  `diff` = sum_k (r_k + 1/r_k) - 4 ~ (r - 1)^2 cancels: plain pointwise relative error is
    p99 9e-4, p99.9 2.3e-3, max 3.8e-2 (at |diff| ~ 1e-6); against max(|diff|, 1e-2) the maximum is 4.4e-3.
    grad tau per component, relative to the norm of its endpoint's gradient: p50 1e-3, p99 8.5e-3, p99.9 2e-2, max 0.27
    (pairs whose gradient at one endpoint nearly vanishes; north_star does not bound grad tau, Speed -- which is built
    from it -- is bounded above)."""
    from oracle import pntf_oracle as O
    P = O.init_params(0)
    B, c, s = _workload(d, E, N, margin)
    variant = capi.LOSS_SINGLE if single else capi.LOSS_MULTI
    tau, dtau, sp, sp2, diff, loss_n = _cuda_outputs(capi, _flat(P), B, c, s, E, N, d, variant, capi.PREC_TC)
    tau_o, dtau_o, sp_o, sp2_o, diff_o, loss_o = _oracle(P, c.cpu(), s.cpu(), B.cpu(), E, N, single)
    assert _rel(tau, tau_o).max() < 1e-3, label
    assert _rel(sp, sp_o).max() < 1e-2, label
    assert _rel(sp2, sp2_o).max() < 1e-2, label
    assert abs(loss_n - loss_o) / abs(loss_o) < 1e-3, label          # north_star: 1e-2
    rd = _rel(diff, diff_o)
    assert np.quantile(rd, 0.999) < 1e-2, label
    assert _rel(diff, diff_o, 1e-2).max() < 1e-2, label
    gn = np.linalg.norm(dtau_o.numpy().reshape(-1, 2, d), axis=2).repeat(d, axis=1)
    rg = np.abs(dtau.numpy() - dtau_o.numpy()) / gn
    assert np.quantile(rg, 0.99) < 2e-2 and np.quantile(rg, 0.999) < 5e-2, label
    assert np.abs(dtau.numpy() - dtau_o.numpy()).max() / np.abs(dtau_o.numpy()).max() < 1e-2, label


@pytest.mark.parametrize("d,N,margin,single", [(3, 20000, 0.5 / 5, False), (6, 5000, 0.5 / 12, True)])
def test_fp32_verify_mode_pointwise_1e5(capi, d, N, margin, single):
    """north_star: within rel 1e-5 in the fp32-SIMT verify mode.  tau, Speed, Speed2 and the loss are asserted directly
    and pointwise (measured maxima 2.7e-7, 6.4e-7, -, 1.4e-7).  `diff` cancels ((r - 1)^2 from O(1) terms): the
    reference's own fp32 arithmetic is off by up to 3.6e-4 relative from an fp64 evaluation there, so it is judged with
    the fp64 oracle as arbiter: |cuda - o64| <= |ref32 - o64| + 5e-5 max(|o64|, 1e-2) (measured excess: 2.3e-5 of that
    scale at d = 3, 1.1e-5 at d = 6; the two fp32 evaluations round differently, neither is systematically closer:
    quantiles of |cuda - o64| are at or below those of |ref32 - o64|, p99.9 4.0e-5 vs 1.0e-4 relative at d = 3)."""
    from oracle import pntf_oracle as O
    P = O.init_params(0)
    B, c, s = _workload(d, 1, N, margin)
    variant = capi.LOSS_SINGLE if single else capi.LOSS_MULTI
    tau, dtau, sp, sp2, diff, loss_n = _cuda_outputs(capi, _flat(P), B, c, s, 1, N, d, variant, capi.PREC_FP32)
    tau_o, dtau_o, sp_o, sp2_o, diff_o, loss_o = _oracle(P, c.cpu(), s.cpu(), B.cpu(), 1, N, single)
    P64 = O.init_params(0, dtype=torch.float64)
    tau64, dtau64, sp64, sp264, diff64, loss64 = _oracle(P64, c.cpu(), s.cpu(), B.cpu(), 1, N, single, dtype=torch.float64)
    assert _rel(tau, tau_o).max() < 1e-5
    assert _rel(sp, sp_o).max() < 1e-5
    assert _rel(sp2, sp2_o).max() < 1e-5
    assert abs(loss_n - loss_o) / abs(loss_o) < 1e-5
    assert abs(loss_n - loss64) / abs(loss64) < 1e-5
    scale = np.maximum(np.abs(diff64.numpy()), 1e-2)
    err_cuda = np.abs(diff.numpy().astype(np.float64) - diff64.numpy())
    err_ref = np.abs(diff_o.numpy().astype(np.float64) - diff64.numpy())
    assert ((err_cuda - err_ref) / scale).max() < 5e-5
    assert np.quantile(err_cuda / scale, 0.999) <= np.quantile(err_ref / scale, 0.999) + 1e-6
    assert (np.abs(dtau.numpy() - dtau_o.numpy()).max(axis=1) / np.abs(dtau_o.numpy()).max(axis=1)).max() < 2e-4


@pytest.mark.parametrize("d,n_train,n_cmp", [(3, 20000, 400), (6, 12000, 200)])
def test_tc_backward_after_200_optimizer_steps(capi, d, n_train, n_cmp):
    """200 AdamW steps in tensor-core mode with alpha ramping 0.5 -> 1 and beta = 1 / loss_n of the previous step (the
    trainer's rules, M:1011,1109, compressed into 200 steps), then: the loss is finite throughout; no operand or adjoint
    tile saturates fp16 (largest magnitudes stay 8x below 65 504 -- before the second-order columns were scaled
    (tc_layers.cuh: kLapScale) X_13 reached 6.5e4 and the loss became inf at step 158 of exactly this run); fewer than
    half of the adjoints are subnormal; TC parameter gradients agree with fp32 mode and oracle autograd.

    The 200 steps run with deterministic gradient accumulation (pntf_set_deterministic), which makes the trajectory --
    and so the weights the final comparison is made at -- the same in every run (final loss_n 0.138052 at d = 3,
    0.106132 at d = 6; worst per-tensor max-norm error 1.9e-2 / 0.5e-2).  With the default fp32 atomics the last bits of
    the gradients differ from run to run, 200 steps of beta = 1 / loss_n amplify that into different trajectories
    (final loss_n 0.11 ... 0.33 over three runs, scripts/probe_200step.py) and the error measured on a 200-400 pair
    slice at whatever weights a run ends at ranged from 0.5e-2 to 6e-2: a property of the end point, not of the kernel,
    and nothing a fixed bound can be asserted on."""
    from oracle import pntf_oracle as O
    from pntf import synth
    from pntf.model import Model_multi
    torch.manual_seed(0)
    m = Model_multi("", "", d, 1, device=DEV)
    m.init_network()
    m.network.precision = "tc"
    fp, opt = m.network.flat(), m.optimizer
    B = synth.sample_B(d, DEV, seed=1, E=1).contiguous()
    c, s = synth.sample_pairs(n_train, d, DEV, seed=2, margin=0.5 / 12, near_obstacle=True)
    beta, hist = 1.0, []
    capi.set_deterministic(True)
    try:
        for it in range(200):
            alpha = 0.5 + 0.5 * it / 199
            fp.grad.zero_()
            lo, _ = capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, 1, n_train, d, alpha, beta, GAMMA, grad_out=fp.grad,
                              prec=capi.PREC_TC)
            ln = float(lo[1])
            assert np.isfinite(ln) and ln > 0, (it, ln)
            hist.append(ln)
            opt.step()
            beta = 1.0 / ln
    finally:
        capi.set_deterministic(False)
    assert min(hist[100:]) < 0.5 * hist[0]
    # fp16 ranges at the full training batch
    g_full = torch.zeros_like(fp.flat)
    capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, 1, n_train, d, 1.0, beta, GAMMA, grad_out=g_full, prec=capi.PREC_TC)
    st = capi.adjoint_stats(1, n_train, d)
    assert st["inf_nan"] == 0 and st["max_abs"] < 65504 / 8, st
    assert st["subnormal"] < 0.5 * st["nonzero"], st
    for l in list(range(14)) + [100]:    # 100 + 100 = layer 200: the encoder outputs Y5
        try:
            sx = capi.adjoint_stats(1, n_train, d, layer=100 + l)
        except RuntimeError as e:        # a -DPNTF_STASH_REGEN build regenerates X_0 / X_6 instead of stashing them
            assert l in (0, 6) and "not stashed" in str(e), (l, e)
            continue
        assert sx["inf_nan"] == 0 and sx["max_abs"] < 65504 / 8, (l, sx)
    assert bool(torch.isfinite(g_full).all())
    # gradients on a slice the oracle can differentiate
    cc, ss = c[:n_cmp].contiguous(), s[:n_cmp].contiguous()
    g_tc, g_32 = torch.zeros_like(fp.flat), torch.zeros_like(fp.flat)
    capi.loss(capi.LOSS_MULTI, fp.flat, B, cc, ss, 1, n_cmp, d, 1.0, beta, GAMMA, grad_out=g_tc, prec=capi.PREC_TC)
    capi.loss(capi.LOSS_MULTI, fp.flat, B, cc, ss, 1, n_cmp, d, 1.0, beta, GAMMA, grad_out=g_32, prec=capi.PREC_FP32)
    P = {k: v.detach().cpu().clone() for k, v in fp.views.items()}
    _, _, _, grads = O.loss_and_grads(P, cc.cpu().unsqueeze(0), ss.cpu().unsqueeze(0), B.cpu(), beta, GAMMA, "multi")
    worst_tc, worst_l2, worst_32 = 0.0, 0.0, 0.0
    for k in P:
        if grads[k] is None:
            continue
        o, n = fp.offsets[k], P[k].numel()
        go = grads[k].reshape(-1).double()
        a, b = g_tc[o:o + n].cpu().double(), g_32[o:o + n].cpu().double()
        worst_tc = max(worst_tc, float((a - go).abs().max() / go.abs().max()))
        worst_l2 = max(worst_l2, float(torch.linalg.norm(a - go) / torch.linalg.norm(go)))
        worst_32 = max(worst_32, float((b - go).abs().max() / go.abs().max()))
    assert worst_32 < 2e-3, worst_32
    assert worst_tc < 3e-2 and worst_l2 < 3e-2, (worst_tc, worst_l2)
