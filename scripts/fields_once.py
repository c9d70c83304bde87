"""One launch each of the value-only (TravelTimes) and first-order (Speed) forward at 1 Mi pairs, for ncu."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from pntf.model import Model_multi
dev = torch.device("cuda:0")
d, n = 3, 1 << 20
torch.manual_seed(0)
m = Model_multi("", "", d, 1, device=str(dev)); m.init_network()
flat = m.network.flat().flat
B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
coords, _ = synth.sample_pairs(n, d, dev, seed=2, margin=0.5 / 12)
for kind in (capi.FIELD_TRAVEL_TIME, capi.FIELD_SPEED):
    capi.field(kind, flat, B, coords, 1, n, d, prec=capi.PREC_TC)
torch.cuda.synchronize()
print("ok")
