"""Small invocations of every tensor-core and fp32 entry point, as a quick all-entry-points smoke (compute-sanitizer is closed on this pool)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from oracle import pntf_oracle as O
dev = torch.device("cuda:0")
P = O.init_params(0)
flat = torch.cat([P[k].reshape(-1) for k in P]).to(dev).contiguous()
for d in (3, 6):
    E, N = 2, 301
    B = synth.sample_B(d, dev, seed=1, E=E).contiguous()
    c, s = synth.sample_pairs(E * N, d, dev, seed=2)
    for prec in (capi.PREC_TC, capi.PREC_FP32):
        g = torch.zeros_like(flat)
        capi.loss(capi.LOSS_MULTI, flat, B, c, s, E, N, d, 1.0, 1.0, 1e-3, grad_out=g, prec=prec, budget=40 << 20)
        capi.loss(capi.LOSS_SINGLE, flat, B[:1], c[:N], s[:N], 1, N, d, 1.0, 1.0, 1e-3, prec=prec)
        table = torch.cat([c, s], dim=1).contiguous()
        idx = torch.randint(0, E * N, (E * N,), device=dev, dtype=torch.int32)
        capi.loss_gather(capi.LOSS_MULTI, flat, B, table, idx, E, N, d, 1.0, 1.0, 1e-3, grad_out=g, prec=prec, budget=40 << 20)
        for mode in (capi.MODE_VALUE, capi.MODE_GRAD, capi.MODE_LAPLACE):
            capi.forward(mode, flat, B, c, E, N, d, prec=prec)
        for kind in range(4):
            capi.field(kind, flat, B[:1], c[:N], 1, N, d, gamma=1e-3, prec=prec)
        capi.plan(flat, B[:1], synth.sample_queries(77, d, dev), d, 0.03, 0.06, 12, prec=prec)
        torch.cuda.synchronize()
        print("d", d, "prec", prec, "ok", float(g.abs().max()))
