"""Time individual entry points on the GPU (CUDA events): loss forward only (no stash), forward modes, training step."""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from oracle import pntf_oracle as O
dev = torch.device("cuda:0")
d = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
P = O.init_params(0)
flat = torch.cat([P[k].reshape(-1) for k in P]).to(dev).contiguous()
B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
coords, speed = synth.sample_pairs(n, d, dev, seed=2, margin=0.5 / 12)
def timeit(f, reps=3):
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
fl = {3: 14290176, 6: 27269376}[d]
t = timeit(lambda: capi.loss(capi.LOSS_MULTI, flat, B, coords, speed, 1, n, d, 1.0, 1.0, 1e-3, prec=capi.PREC_TC))
print("loss fwd only (TC, no stash): %.2f ms  %.1f TFLOP/s alg" % (t, fl * n / t / 1e9))
t = timeit(lambda: capi.forward(capi.MODE_GRAD, flat, B, coords, 1, n, d, prec=capi.PREC_TC))
print("out_grad (TC): %.2f ms" % t)
grad = torch.zeros_like(flat)
capi.profile_begin()
t = timeit(lambda: capi.loss(capi.LOSS_MULTI, flat, B, coords, speed, 1, n, d, 1.0, 1.0, 1e-3, grad_out=grad, prec=capi.PREC_TC))
prof = capi.profile_end()
print("loss fwd+bwd (TC): %.2f ms  %.1f TFLOP/s alg" % (t, 3 * fl * n / t / 1e9), {k: round(v[0] / 4, 2) for k, v in prof.items() if v[1]})
