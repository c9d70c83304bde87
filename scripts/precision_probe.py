"""CPU probe: effect of rounding MMA operands (activations, weights, adjoints) to fp16/bf16 with fp32 accumulation on
tau / speed / loss / parameter gradients, using the oracle.  Run: python scripts/precision_probe.py"""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pntf_oracle as O

MAXABS = {}
def make_linear(dt, gscale=1.0):
    def q(x): return x.to(dt).to(torch.float32)
    class Lin(torch.autograd.Function):
        @staticmethod
        def forward(ctx, t, w, b):
            tq, wq = q(t), q(w)
            ctx.save_for_backward(tq, wq)
            MAXABS['x'] = max(MAXABS.get('x', 0), float(t.abs().max()))
            y = tq @ wq.T
            y[:, 0, :] = y[:, 0, :] + b
            return y
        @staticmethod
        def backward(ctx, gy):
            tq, wq = ctx.saved_tensors
            g = gy * gscale
            MAXABS['g'] = max(MAXABS.get('g', 0), float(g.abs().max()))
            nz = g[g != 0].abs()
            if nz.numel(): MAXABS['gmin'] = min(MAXABS.get('gmin', 1e30), float(nz.median()))
            gq = q(g)
            gx = (gq @ wq) / gscale
            gw = torch.einsum('rcm,rck->mk', gq, tq) / gscale
            gb = gy[:, 0, :].sum(0)
            return gx, gw, gb
    return lambda t, w, b: Lin.apply(t, w, b)

def run(d, N, dt, gscale):
    rs = np.random.RandomState(3)
    P = O.init_params(1)
    Bw = torch.from_numpy((0.5 * rs.normal(size=(1, d, 128))).astype(np.float32))
    c = torch.from_numpy(rs.uniform(-0.5, 0.5, size=(1, N, 2 * d)).astype(np.float16).astype(np.float32))
    s = torch.from_numpy(rs.uniform(0.1, 1.0, size=(1, N, 2)).astype(np.float32))
    ref = O.loss_and_grads(P, c, s, Bw, 1.0, 1e-3, "multi")
    with torch.no_grad():
        rt, rd, rl = O.forward(P, c[0], Bw[0], 2)
        rsp = O.speed_pred(P, c[0], Bw[0])
    orig = O._linear_cols
    O._linear_cols = make_linear(dt, gscale)
    try:
        MAXABS.clear()
        got = O.loss_and_grads(P, c, s, Bw, 1.0, 1e-3, "multi")
        with torch.no_grad():
            tt, td, tl = O.forward(P, c[0], Bw[0], 2)
            tsp = O.speed_pred(P, c[0], Bw[0])
    finally:
        O._linear_cols = orig
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
    relp = lambda a, b: float(((a - b).abs() / b.abs()).max())
    gerr = max(rel(got[3][k], ref[3][k]) for k in ref[3] if ref[3][k] is not None)
    print("d=%d %s: tau relmax %.2e  dtau %.2e  ltau %.2e  speed relmax(pointwise) %.2e  loss %.2e  diff %.2e  grad %.2e | max|x| %.3g max|g| %.3g med|g| %.3g"
          % (d, str(dt).split('.')[-1], relp(tt, rt), rel(td, rd), rel(tl, rl), relp(tsp, rsp), abs(got[0] - ref[0]) / abs(ref[0]), rel(got[2], ref[2]), gerr,
             MAXABS.get('x', 0), MAXABS.get('g', 0), MAXABS.get('gmin', 0)))

for d in (3, 6):
    for dt in (torch.float16, torch.bfloat16):
        run(d, 1000, dt, 1000.0)
