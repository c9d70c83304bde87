"""Print error metrics of the tensor-core forward against the oracle (run on the GPU box)."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from oracle import pntf_oracle as O
from pntf import capi
DEV = "cuda:0"
def cu(a): return torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float32))).to(DEV)
def rel(a, b): return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / (np.abs(np.asarray(b)).max() + 1e-30))
for d, N in ((3, 5), (3, 100), (6, 50)):
    rs = np.random.RandomState(1)
    P = O.init_params(21)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    Bw = (0.5 * rs.normal(size=(d, 128))).astype(np.float32)
    c = rs.uniform(-0.5, 0.5, size=(N, 2 * d)).astype(np.float16).astype(np.float32)
    with torch.no_grad():
        tau, dtau, ltau = O.forward(P, torch.from_numpy(c), torch.from_numpy(Bw), 2)
    for mode in (capi.MODE_LAPLACE, capi.MODE_GRAD):
        try:
            t2, d2, l2 = capi.forward(mode, flat, cu(Bw).view(1, d, 128), cu(c), 1, N, d, prec=capi.PREC_TC)
            torch.cuda.synchronize()
        except Exception as e:
            print("d=%d N=%d mode=%d FAILED: %s" % (d, N, mode, e)); continue
        print("d=%d N=%d mode=%d: tau %.3e dtau %.3e ltau %s" % (d, N, mode, rel(t2.cpu(), tau[:, 0]), rel(d2.cpu(), dtau),
              "%.3e" % rel(l2.cpu(), ltau) if l2 is not None else "-"))
        if rel(t2.cpu(), tau[:, 0]) > 1e-2:
            print("  tau gpu", t2.cpu().numpy()[:8]); print("  tau ref", tau[:8, 0].numpy())
            print("  dtau gpu", d2.cpu().numpy()[0]); print("  dtau ref", dtau[0].numpy())

from pntf.params import tensor_offsets, tensor_shapes
offs, shapes = tensor_offsets(), tensor_shapes()
for d, E, N, variant in ((3, 1, 40, "multi"), (3, 2, 1500, "multi"), (6, 1, 100, "single")):
    rs = np.random.RandomState(3)
    P = O.init_params(5)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    Bw = (0.5 * rs.normal(size=(E, d, 128))).astype(np.float32)
    c = rs.uniform(-0.5, 0.5, size=(E, N, 2 * d)).astype(np.float16).astype(np.float32)
    s = rs.uniform(0.1, 1.0, size=(E, N, 2)).astype(np.float32)
    if variant == "multi":
        loss, loss_n, diff, grads = O.loss_and_grads(P, torch.from_numpy(c), torch.from_numpy(s), torch.from_numpy(Bw), 1.0, 1e-3, "multi")
    else:
        loss, loss_n, diff, grads = O.loss_and_grads(P, torch.from_numpy(c[0]), torch.from_numpy(s[0]), torch.from_numpy(Bw[0]), 1.0, 1e-3, "single")
    grad = torch.zeros_like(flat)
    v = capi.LOSS_MULTI if variant == "multi" else capi.LOSS_SINGLE
    try:
        lo, df = capi.loss(v, flat, cu(Bw), cu(c).view(E * N, 2 * d), cu(s).view(E * N, 2), E, N, d, 1.0, 1.0, 1e-3, grad_out=grad, prec=capi.PREC_TC)
        torch.cuda.synchronize()
    except Exception as e:
        print("train d=%d N=%d FAILED: %s" % (d, N, e)); continue
    print("train d=%d E=%d N=%d: loss gpu %.6e ref %.6e ; diff err %.2e" % (d, E, N, lo[0].item(), loss.item(), rel(df.cpu(), diff.reshape(-1))))
    g = grad.cpu()
    for k, ref in grads.items():
        n = int(np.prod(shapes[k]))
        gk = g[offs[k]:offs[k] + n].view(shapes[k])
        if ref is None:
            print("   %-22s zero=%s" % (k, bool((gk == 0).all())))
        else:
            print("   %-22s rel err %.3e   (|ref|max %.3e, |gpu|max %.3e)" % (k, rel(gk, ref), float(ref.abs().max()), float(gk.abs().max())))
