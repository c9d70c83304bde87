"""Spread of tests/test_gpu_north_star.py::test_tc_backward_after_200_optimizer_steps's gradient errors over repeated runs
(default atomics vs deterministic accumulation) and over the size of the comparison slice."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "p-ntfields_b200")]
import numpy as np, torch
from pntf import capi, synth
from pntf.model import Model_multi
from oracle import pntf_oracle as O
DEV = "cuda:0"; GAMMA = 1e-3

def run(d, n_train, det, cmps):
    capi.set_deterministic(det)
    torch.manual_seed(0)
    m = Model_multi("", "", d, 1, device=DEV); m.init_network(); m.network.precision = "tc"
    fp, opt = m.network.flat(), m.optimizer
    B = synth.sample_B(d, DEV, seed=1, E=1).contiguous()
    c, s = synth.sample_pairs(n_train, d, DEV, seed=2, margin=0.5 / 12, near_obstacle=True)
    beta = 1.0
    for it in range(200):
        alpha = 0.5 + 0.5 * it / 199
        fp.grad.zero_()
        lo, _ = capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, 1, n_train, d, alpha, beta, GAMMA, grad_out=fp.grad, prec=capi.PREC_TC)
        ln = float(lo[1]); opt.step(); beta = 1.0 / ln
    capi.set_deterministic(False)
    out = []
    for n_cmp in cmps:
        cc, ss = c[:n_cmp].contiguous(), s[:n_cmp].contiguous()
        g_tc, g_32 = torch.zeros_like(fp.flat), torch.zeros_like(fp.flat)
        capi.loss(capi.LOSS_MULTI, fp.flat, B, cc, ss, 1, n_cmp, d, 1.0, beta, GAMMA, grad_out=g_tc, prec=capi.PREC_TC)
        capi.loss(capi.LOSS_MULTI, fp.flat, B, cc, ss, 1, n_cmp, d, 1.0, beta, GAMMA, grad_out=g_32, prec=capi.PREC_FP32)
        # fp32 mode as the reference here (agrees with oracle autograd to 2e-3, asserted in the test)
        worst = (0, None); worst_l2 = 0; tot_num = 0; tot_den = 0
        for k, o in fp.offsets.items():
            n = fp.views[k].numel()
            a, b = g_tc[o:o + n].double(), g_32[o:o + n].double()
            if float(b.abs().max()) == 0: continue
            e = float((a - b).abs().max() / b.abs().max())
            if e > worst[0]: worst = (e, k)
            worst_l2 = max(worst_l2, float(torch.linalg.norm(a - b) / torch.linalg.norm(b)))
            tot_num += float(((a - b) ** 2).sum()); tot_den += float((b ** 2).sum())
        out.append((n_cmp, round(worst[0], 4), worst[1], round(worst_l2, 4), round((tot_num / tot_den) ** 0.5, 4)))
    return round(ln, 6), out

for d, n_train in [(6, 12000), (3, 20000)]:
    for det in (True, True, False, False, False):
        print(d, "det" if det else "atomic", run(d, n_train, det, [200, 400, 2000] if d == 6 else [400, 2000]), flush=True)
