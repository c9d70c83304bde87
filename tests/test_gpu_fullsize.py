"""Full-size GPU checks (BASELINE.json configs: 1 Mi point pairs per step, 65 536 path queries) through
size-independent properties, where the CPU oracle would need hours:
  * the travel-time network is symmetric in (start, goal) by construction (S:724-787): swapping the endpoints
    leaves tau unchanged and swaps the two halves of grad tau / diag-Laplacian;
  * the per-pair residuals of a batch do not depend on how the batch is cut, and parameter gradients add up;
  * the tensor-core mode agrees with the fp32 verify mode within north_star's tolerances (tau rel 1e-3, loss rel 1e-2);
  * path integration: step counts within the cap, every taken step has the prescribed length, terminated queries met.
"""
import numpy as np
import pytest
import torch

from oracle import pntf_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
FULL = 1 << 20


@pytest.fixture(scope="module")
def capi():
    from pntf import capi as c
    c.lib()
    return c


@pytest.fixture(scope="module")
def setup():
    from pntf import synth
    P = O.init_params(3)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    out = {}
    for d in (3, 6):
        B = synth.sample_B(d, DEV, seed=1, E=1).contiguous()
        n = FULL if d == 3 else FULL // 4
        coords, speed = synth.sample_pairs(n, d, DEV, seed=5, margin=0.5 / 12)
        out[d] = (B, coords, speed)
    return flat, out


@pytest.mark.parametrize("d", [3, 6])
def test_swap_symmetry_full_size(capi, setup, d):
    flat, data = setup
    B, c, _ = data[d]
    n = c.shape[0]
    sw = torch.cat([c[:, d:], c[:, :d]], dim=1).contiguous()
    t1, g1, l1 = capi.forward(capi.MODE_LAPLACE, flat, B, c, 1, n, d, prec=capi.PREC_TC)
    t2, g2, l2 = capi.forward(capi.MODE_LAPLACE, flat, B, sw, 1, n, d, prec=capi.PREC_TC)
    torch.cuda.synchronize()
    assert torch.isfinite(t1).all() and torch.isfinite(g1).all() and torch.isfinite(l1).all()
    assert float(((t1 - t2).abs() / t1.abs()).max()) < 1e-5
    # the two orderings evaluate the combine gates as s and 1 - s: equal up to rounding of the gate (most rows are
    # bit-identical); the bound is a tenth of the tensor-core mode's derivative tolerance
    scale = float(g1.abs().max())
    assert float((g1[:, :d] - g2[:, d:]).abs().max()) < 1e-3 * scale
    assert float((g1[:, d:] - g2[:, :d]).abs().max()) < 1e-3 * scale
    lscale = float(l1.abs().max())
    assert float((l1[:, :d] - l2[:, d:]).abs().max()) < 3e-3 * lscale


@pytest.mark.parametrize("prec_name", ["tc", "fp32"])
def test_batch_cut_independence_and_gradient_additivity(capi, setup, prec_name):
    flat, data = setup
    d = 3
    B, c, s = data[d]
    n = FULL if prec_name == "tc" else FULL // 8
    c, s = c[:n].contiguous(), s[:n].contiguous()
    prec = capi.PREC_TC if prec_name == "tc" else capi.PREC_FP32
    g_full = torch.zeros_like(flat)
    lo, diff = capi.loss(capi.LOSS_MULTI, flat, B, c, s, 1, n, d, 1.0, 1.0, 1e-3, grad_out=g_full, prec=prec)
    parts = 4
    m = n // parts
    g_sum = torch.zeros_like(flat)
    diffs = []
    for k in range(parts):
        g = torch.zeros_like(flat)
        _, dk = capi.loss(capi.LOSS_MULTI, flat, B, c[k * m:(k + 1) * m].contiguous(), s[k * m:(k + 1) * m].contiguous(),
                          1, m, d, 1.0, 1.0, 1e-3, grad_out=g, prec=prec)
        diffs.append(dk.clone())
        g_sum += g / parts
    torch.cuda.synchronize()
    assert torch.isfinite(diff).all()
    assert torch.equal(torch.cat(diffs), diff)                      # per-pair residuals: bit-identical
    # loss_n = mean(diff) + 0.01 |B|^2 / (E N)   (M:947)
    want = diff.double().mean().item() + 0.01 * float((B.double() ** 2).sum()) / n
    assert abs(lo[1].item() - want) / want < 1e-5
    assert float((g_full - g_sum).abs().max()) < 2e-4 * float(g_full.abs().max())


def test_tc_agrees_with_fp32_mode_at_scale(capi, setup):
    flat, data = setup
    d = 3
    B, c, s = data[d]
    n = FULL // 4
    c, s = c[:n].contiguous(), s[:n].contiguous()
    t_tc, g_tc, _ = capi.forward(capi.MODE_LAPLACE, flat, B, c, 1, n, d, prec=capi.PREC_TC)
    t_32, g_32, _ = capi.forward(capi.MODE_LAPLACE, flat, B, c, 1, n, d, prec=capi.PREC_FP32)
    lo_tc, _ = capi.loss(capi.LOSS_MULTI, flat, B, c, s, 1, n, d, 1.0, 1.0, 1e-3, prec=capi.PREC_TC)
    lo_32, _ = capi.loss(capi.LOSS_MULTI, flat, B, c, s, 1, n, d, 1.0, 1.0, 1e-3, prec=capi.PREC_FP32)
    torch.cuda.synchronize()
    assert float(((t_tc - t_32).abs() / t_32.abs()).max()) < 1e-3          # tau within rel 1e-3 (north_star)
    assert float((g_tc - g_32).abs().max()) < 1e-2 * float(g_32.abs().max())
    assert abs(lo_tc[1].item() - lo_32[1].item()) / abs(lo_32[1].item()) < 1e-2   # loss within rel 1e-2


@pytest.mark.parametrize("d,eta,thr,cap", [(3, 0.03, 0.06, 500), (6, 0.015, 0.03, 300)])
def test_plan_invariants_65536_queries(capi, setup, d, eta, thr, cap):
    from pntf import synth
    flat, data = setup
    B = data[d][0]
    Q = 65536
    q = synth.sample_queries(Q, d, DEV)
    way, nst = capi.plan(flat, B, q, d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC)
    torch.cuda.synchronize()
    assert torch.isfinite(way).all()
    assert int(nst.min()) >= 0 and int(nst.max()) <= cap + 1
    assert torch.equal(way[:, 0], q)                                     # row 0 is the query
    idx = torch.arange(Q, device=DEV)
    last = way[idx, nst.long()]
    gap = (last[:, d:] - last[:, :d]).norm(dim=1)
    done = nst <= cap
    assert bool((gap[done] <= thr * (1 + 1e-5)).all())                   # terminated queries met
    # Gradient returns v / |v|^2 per endpoint (S:1270-1282): a step moves each endpoint by eta / |v|; it must be
    # finite and positive for every step taken
    step1 = (way[:, 1] - way[:, 0])
    moved = nst >= 1
    assert bool((step1[moved].norm(dim=1) > 0).all())
    # rows after the last step repeat the final position
    tail = way[idx, torch.clamp(nst.long() + 1, max=cap + 1)]
    assert torch.equal(tail, last)


@pytest.mark.parametrize("d", [3, 6])
def test_value_only_rows_equal_the_value_column_of_the_first_order_rows(capi, setup, d):
    """NN.out / Model.TravelTimes run with one tile row per endpoint (64 pairs per tile); the value column of a pair
    does not depend on which other rows share its tile, so tau must equal the first-order kernel's bit for bit, also
    for a ragged count, and TravelTimes = |x1 - x0| / tau (S:1169-1188)."""
    flat, data = setup
    B, c, _ = data[d]
    for n in (c.shape[0], 64 * 148 * 3 + 29):
        cc = c[:n].contiguous()
        t0, _, _ = capi.forward(capi.MODE_VALUE, flat, B, cc, 1, n, d, prec=capi.PREC_TC)
        t1, _, _ = capi.forward(capi.MODE_GRAD, flat, B, cc, 1, n, d, prec=capi.PREC_TC)
        tt = capi.field(capi.FIELD_TRAVEL_TIME, flat, B, cc, 1, n, d, prec=capi.PREC_TC)
        torch.cuda.synchronize()
        assert torch.isfinite(t0).all()
        assert torch.equal(t0, t1)
        want = (cc[:, d:] - cc[:, :d]).norm(dim=1) / t0
        assert float(((tt.view(-1) - want).abs() / want.abs().clamp_min(1e-20)).max()) < 1e-5


@pytest.mark.parametrize("d,eta,thr,cap", [(3, 0.03, 0.06, 20), (6, 0.015, 0.03, 45)])
def test_plan_does_not_depend_on_slot_scheduling(capi, setup, d, eta, thr, cap):
    """Finished queries hand their tile slot to the next unstarted query (dynamic refill), so which queries share a
    tile differs from run to run and with the batch; a query's rows are independent of its tile mates, so its
    waypoints must be bit-identical when it is planned alone in a small batch, in a permuted batch, or among 65 536;
    the low cap also exercises the step-cap exit next to the met exit."""
    from pntf import synth
    flat, data = setup
    B = data[d][0]
    Q = 65536
    q = synth.sample_queries(Q, d, DEV)
    q[5, d:] = q[5, :d] + 0.1 * thr                                      # already met: zero steps
    way, nst = capi.plan(flat, B, q, d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC)
    sub = torch.arange(0, Q, 97, device=DEV)
    way_s, nst_s = capi.plan(flat, B, q[sub].contiguous(), d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC)
    perm = torch.randperm(Q, device=DEV, generator=torch.Generator(device=DEV).manual_seed(3))
    way_p, nst_p = capi.plan(flat, B, q[perm].contiguous(), d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC)
    torch.cuda.synchronize()
    assert int(nst[5]) == 0 and torch.equal(way[5], q[5].expand(cap + 2, 2 * d))
    assert int(nst.max()) == cap + 1 and int((nst <= cap).sum()) > 0     # both exits taken
    assert torch.equal(nst_s, nst[sub]) and torch.equal(way_s, way[sub])
    assert torch.equal(nst_p, nst[perm]) and torch.equal(way_p, way[perm])
