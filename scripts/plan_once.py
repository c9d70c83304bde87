"""Times the 65 536-query path integration (BASELINE configs[4]) at several query counts: python scripts/plan_once.py"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from pntf.model import Model_multi
dev = torch.device("cuda:0")
for d, eta, thr, cap in ((3, 0.03, 0.06, 500), (6, 0.015, 0.03, 300)):
    torch.manual_seed(0)
    m = Model_multi("", "", d, 1, device=str(dev)); m.init_network()
    flat = m.network.flat().flat
    B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
    for Q in (65536, 8192):
        q = synth.sample_queries(65536, d, dev)[:Q].contiguous()
        for fill in (True, False):
            capi.plan(flat, B, q, d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC, fill_tail=fill)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            way, nst = capi.plan(flat, B, q, d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC, fill_tail=fill)
            b.record(); torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            print("d=%d Q=%d fill=%d: %.3f ms, %.3f M paths/s, mean steps %.2f max %d" % (d, Q, fill, ms, Q / ms / 1e3, float(nst.float().mean()), int(nst.max())), flush=True)
