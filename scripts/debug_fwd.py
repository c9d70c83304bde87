import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from oracle import pntf_oracle as O
from pntf import capi
dev = "cuda:0"
def T(a): return torch.from_numpy(np.ascontiguousarray(a))
for d in (3, 6):
    for E, N in ((1, 24), (2, 24), (1, 50)):
        rs = np.random.RandomState(5)
        P = O.init_params(3)
        flat = torch.cat([P[k].reshape(-1) for k in P]).to(dev)
        Bw = (0.5 * rs.normal(size=(E, d, 128))).astype(np.float32)
        c = rs.uniform(-0.5, 0.5, size=(E, N, 2 * d)).astype(np.float32)
        s = rs.uniform(0.1, 1.0, size=(E, N, 2)).astype(np.float32)
        for mode in (0, 1, 2):
            tau, dtau, ltau = capi.forward(mode, flat, T(Bw).to(dev), T(c).view(E * N, 2 * d).to(dev), E, N, d)
            torch.cuda.synchronize()
            for e in range(E):
                with torch.no_grad():
                    rt, rd, rl = O.forward(P, T(c[e]), T(Bw[e]), mode)
                sl = slice(e * N, (e + 1) * N)
                msg = "d=%d E=%d N=%d mode=%d env=%d: tau %.2e" % (d, E, N, mode, e, (tau[sl].cpu() - rt[:, 0]).abs().max())
                if mode >= 1:
                    er = (dtau[sl].cpu() - rd).abs().max(dim=1).values / rd.abs().max()
                    msg += " dtau %.2e (bad rows %s)" % (er.max(), (er > 1e-4).nonzero().flatten().tolist()[:12])
                if mode >= 2:
                    er = (ltau[sl].cpu() - rl).abs().max(dim=1).values / rl.abs().max()
                    msg += " ltau %.2e (bad rows %s)" % (er.max(), (er > 1e-3).nonzero().flatten().tolist()[:12])
                print(msg)
        # loss fwd and loss bwd
        lo, diff = capi.loss(capi.LOSS_MULTI, flat, T(Bw).to(dev), T(c).view(E * N, 2 * d).to(dev), T(s).view(E * N, 2).to(dev), E, N, d, 1.0, 1.0, 1e-3)
        grad = torch.zeros_like(flat)
        lo2, diff2 = capi.loss(capi.LOSS_MULTI, flat, T(Bw).to(dev), T(c).view(E * N, 2 * d).to(dev), T(s).view(E * N, 2).to(dev), E, N, d, 1.0, 1.0, 1e-3, grad_out=grad)
        torch.cuda.synchronize()
        loss, loss_n, rdiff, grads = O.loss_and_grads(P, T(c), T(s), T(Bw), 1.0, 1e-3, "multi")
        print("  loss fwd %.6e bwd %.6e oracle %.6e ; diff err fwd %.2e bwd %.2e" % (lo[0].item(), lo2[0].item(), loss.item(),
              (diff.cpu() - rdiff.reshape(-1)).abs().max(), (diff2.cpu() - rdiff.reshape(-1)).abs().max()))
        g = torch.cat([(grads[k] if grads[k] is not None else torch.zeros_like(P[k])).reshape(-1) for k in P])
        print("  grad rel err %.2e" % ((grad.cpu() - g).abs().max() / g.abs().max()))
