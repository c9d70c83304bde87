"""PNTF_TRACE=1 python scripts/trace_fwd.py : one training-forward and one no-stash forward launch with time stamps."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from oracle import pntf_oracle as O
dev = torch.device("cuda:0")
d, n = 3, 1 << 16
P = O.init_params(0)
flat = torch.cat([P[k].reshape(-1) for k in P]).to(dev).contiguous()
B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
coords, speed = synth.sample_pairs(n, d, dev, seed=2, margin=0.5 / 12)
print("=== no stash", file=sys.stderr)
capi.loss(capi.LOSS_MULTI, flat, B, coords, speed, 1, n, d, 1.0, 1.0, 1e-3, prec=capi.PREC_TC)
torch.cuda.synchronize()
print("=== training", file=sys.stderr)
grad = torch.zeros_like(flat)
capi.loss(capi.LOSS_MULTI, flat, B, coords, speed, 1, n, d, 1.0, 1.0, 1e-3, grad_out=grad, prec=capi.PREC_TC)
torch.cuda.synchronize()
