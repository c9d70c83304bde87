"""TC plan vs fp32 SIMT plan vs oracle (run on the GPU box)."""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from oracle import pntf_oracle as O
from pntf import capi
DEV = "cuda:0"
def cu(a): return torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float32))).to(DEV)
for d, Q, cap, mode in ((3, 19, 12, "exact"), (3, 300, 40, "compat"), (6, 50, 20, "exact")):
    rs = np.random.RandomState(9)
    P = O.init_params(8)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    Bw = (0.5 * rs.normal(size=(d, 128))).astype(np.float32)
    sg = rs.uniform(-0.4, 0.4, size=(Q, 2 * d)).astype(np.float32)
    eta, thr = (0.03, 0.06) if d == 3 else (0.015, 0.03)
    fl = capi.GRAD_COMPAT if mode == "compat" else capi.GRAD_EXACT
    w1, n1 = capi.plan(flat, cu(Bw).view(1, d, 128), cu(sg), d, eta, thr, cap, fl, prec=capi.PREC_FP32)
    w2, n2 = capi.plan(flat, cu(Bw).view(1, d, 128), cu(sg), d, eta, thr, cap, fl, prec=capi.PREC_TC)
    torch.cuda.synchronize()
    n1, n2, w1, w2 = n1.cpu().numpy(), n2.cpu().numpy(), w1.cpu().numpy(), w2.cpu().numpy()
    same = n1 == n2
    print("d=%d Q=%d %s: nsteps equal %d/%d, mean steps %.2f, max|dw| (equal-step queries) %.3e, first-step max|dw| %.3e"
          % (d, Q, mode, same.sum(), Q, n1.mean(), np.abs(w1[same] - w2[same]).max(), np.abs(w1[:, 1] - w2[:, 1]).max()))
    if (~same).any():
        q = np.where(~same)[0][0]
        print("  mismatch q=%d n=%d/%d" % (q, n1[q], n2[q]))
# throughput
for d in (3, 6):
    Q = 65536
    rs = np.random.RandomState(3)
    P = O.init_params(0)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    Bw = cu((0.5 * rs.normal(size=(1, d, 128))).astype(np.float32))
    sg = cu(rs.uniform(-0.4, 0.4, size=(Q, 2 * d)).astype(np.float32))
    eta, thr, cap = (0.03, 0.06, 500) if d == 3 else (0.015, 0.03, 300)
    for prec in (capi.PREC_FP32, capi.PREC_TC):
        capi.plan(flat, Bw, sg[:2048].contiguous(), d, eta, thr, cap, prec=prec); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); w, n = capi.plan(flat, Bw, sg, d, eta, thr, cap, prec=prec); b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b); st = float(n.float().sum())
        print("d=%d prec=%d: %.2f ms, %.3e paths/s, %.3e path-steps/s, mean steps %.2f" % (d, prec, ms, Q / ms * 1e3, st / ms * 1e3, st / Q))
