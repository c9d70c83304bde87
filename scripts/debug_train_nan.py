"""Why did 200 TC training steps go NaN?  Same loop in fp32 and tc; per-step loss, adjoint stats of the TC backward."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from pntf.model import Model_multi
dev = torch.device("cuda:0")
d, n = 3, 20000
B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
c, s = synth.sample_pairs(n, d, dev, seed=2, margin=0.5 / 12, near_obstacle=True)
for prec in ("tc",):
    torch.manual_seed(0)
    m = Model_multi("", "", d, 1, device=str(dev)); m.init_network(); m.network.precision = prec
    fp, opt = m.network.flat(), m.optimizer
    pv = capi.PREC_TC if prec == "tc" else capi.PREC_FP32
    beta = 1.0
    for it in range(200):
        alpha = 0.5 + 0.5 * it / 199
        fp.grad.zero_()
        lo, diff = capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, 1, n, d, alpha, beta, 1e-3, grad_out=fp.grad, prec=pv)
        ln = float(lo[1])
        gmax = float(fp.grad.abs().max())
        st = capi.adjoint_stats(1, n, d) if prec == "tc" else {}
        if prec == "tc" and (it % 20 == 0 or it > 195):
            st["Xmax"] = [round(capi.adjoint_stats(1, n, d, layer=100 + l)["max_abs"], 1) for l in range(14)]
            st["Xbad"] = [capi.adjoint_stats(1, n, d, layer=100 + l)["inf_nan"] for l in range(14)]
            st["Gmax"] = [round(capi.adjoint_stats(1, n, d, layer=l)["max_abs"], 2) for l in range(13)]
        print(prec, it, "alpha %.3f beta %.3g loss_n %.6g diff_max %.4g grad_max %.4g" % (alpha, beta, ln, float(diff.max()), gmax), st, flush=True)
        if not (ln == ln) or not (gmax == gmax) or gmax > 1e30:
            break
        opt.step()
        beta = 1.0 / ln
