"""Collects the error statistics behind tests/test_gpu_north_star.py (prints one JSON per experiment)."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from oracle import pntf_oracle as O
dev = torch.device("cuda:0")
torch.set_num_threads(os.cpu_count())

def q(a):
    a = np.asarray(a, np.float64)
    return {"p50": float(np.quantile(a, .5)), "p99": float(np.quantile(a, .99)), "p999": float(np.quantile(a, .999)), "max": float(a.max())}

def rel(a, b, floor=0.0):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor)

def oracle_chunks(P, c, s, B, E, N, gamma, dtype=torch.float32, chunk=10000):
    Pd = {k: v.to(dtype) for k, v in P.items()}
    taus, dtaus, speeds, diffs = [], [], [], []
    with torch.no_grad():
        for e in range(E):
            for lo in range(0, N, chunk):
                cc = c[e * N + lo:e * N + min(lo + chunk, N)].to(dtype)
                ss = s[e * N + lo:e * N + min(lo + chunk, N)].to(dtype)
                Bw = B[e].to(dtype)
                tau, dtau, ltau = O.forward(Pd, cc, Bw, 2)
                _, ln, diff = O.loss_multi(Pd, cc.unsqueeze(0), ss.unsqueeze(0), Bw.unsqueeze(0), 1.0, gamma)
                taus.append(tau[:, 0]); dtaus.append(dtau); diffs.append(diff[0])
                speeds.append(O.speed_pred(Pd, cc, Bw))
    return torch.cat(taus), torch.cat(dtaus), torch.cat(speeds), torch.cat(diffs)

def forward_stats(d, E, N, margin, prec, label, seed=0, arb64=False):
    P = O.init_params(seed)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(dev).contiguous()
    B = synth.sample_B(d, dev, seed=1, E=E).contiguous()
    c, s = synth.sample_pairs(E * N, d, dev, seed=2, margin=margin, near_obstacle=True)
    t0 = time.time()
    tau_o, dtau_o, sp_o, diff_o = oracle_chunks(P, c.cpu(), s.cpu(), B.cpu(), E, N, 1e-3)
    t_or = time.time() - t0
    pv = capi.PREC_TC if prec == "tc" else capi.PREC_FP32
    tau, dtau, ltau = capi.forward(capi.MODE_LAPLACE, flat, B, c, E, N, d, prec=pv)
    lo, diff = capi.loss(capi.LOSS_MULTI, flat, B, c, s, E, N, d, 1.0, 1.0, 1e-3, prec=pv)
    sp = torch.cat([capi.field(capi.FIELD_SPEED, flat, B[e:e + 1].contiguous(), c[e * N:(e + 1) * N].contiguous(), 1, N, d, prec=pv) for e in range(E)])
    torch.cuda.synchronize()
    loss_o = float(diff_o.double().sum() / (E * N) + 0.01 * float((B.double() ** 2).sum()) / (E * N))
    gn = np.linalg.norm(dtau_o.numpy().reshape(-1, 2, d), axis=2).repeat(d, axis=1)
    out = {"label": label, "d": d, "E": E, "N": N, "prec": prec, "oracle_s": t_or,
           "speed_min": float(s.min()), "speed_frac_at_clip": float((s <= 0.1 + 1e-6).float().mean()),
           "tau": q(rel(tau.cpu(), tau_o)), "speed": q(rel(sp.cpu(), sp_o)), "diff": q(rel(diff.cpu(), diff_o)),
           "diff_abs_floor1e-2": q(rel(diff.cpu(), diff_o, 1e-2)),
           "diff_min_abs": float(diff_o.abs().min()),
           "dtau_vs_endpoint_norm": q(np.abs(dtau.cpu().numpy() - dtau_o.numpy()) / gn),
           "loss_n_rel": abs(float(lo[1]) - loss_o) / abs(loss_o)}
    if arb64:
        tau64, dtau64, sp64, diff64 = oracle_chunks(P, c.cpu(), s.cpu(), B.cpu(), E, N, 1e-3, torch.float64)
        out["arb"] = {"speed_cuda_vs_64": q(rel(sp.cpu(), sp64)), "speed_ref32_vs_64": q(rel(sp_o, sp64)),
                      "diff_cuda_vs_64": q(rel(diff.cpu(), diff64, 1e-3)), "diff_ref32_vs_64": q(rel(diff_o, diff64, 1e-3)),
                      "tau_cuda_vs_64": q(rel(tau.cpu(), tau64)), "tau_ref32_vs_64": q(rel(tau_o, tau64)),
                      "loss_cuda_vs_64": abs(float(lo[1]) - float(diff64.sum() / (E * N) + 0.01 * float((B.double() ** 2).sum()) / (E * N))),
                      "excess_speed": q(np.maximum(rel(sp.cpu(), sp64) - rel(sp_o, sp64), 0)),
                      "excess_diff": q(np.maximum(rel(diff.cpu(), diff64, 1e-3) - rel(diff_o, diff64, 1e-3), 0))}
    print(json.dumps(out), flush=True)

def trained_grads(d=3, steps=200, n_train=20000, n_cmp=400):
    from pntf.model import Model_multi
    torch.manual_seed(0)
    m = Model_multi("", "", d, 1, device=str(dev)); m.init_network(); m.network.precision = "tc"
    fp, opt = m.network.flat(), m.optimizer
    B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
    c, s = synth.sample_pairs(n_train, d, dev, seed=2, margin=0.5 / 12, near_obstacle=True)
    beta = 1.0
    hist = []
    for it in range(steps):
        alpha = 0.5 + 0.5 * it / (steps - 1)
        fp.grad.zero_()
        lo, _ = capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, 1, n_train, d, alpha, beta, 1e-3, grad_out=fp.grad, prec=capi.PREC_TC)
        opt.step()
        ln = float(lo[1]); hist.append(ln)
        beta = 1.0 / ln
    cc, ss = c[:n_cmp].contiguous(), s[:n_cmp].contiguous()
    g_tc = torch.zeros_like(fp.flat); g_32 = torch.zeros_like(fp.flat)
    capi.loss(capi.LOSS_MULTI, fp.flat, B, cc, ss, 1, n_cmp, d, 1.0, beta, 1e-3, grad_out=g_tc, prec=capi.PREC_TC)
    st = capi.adjoint_stats(1, n_cmp, d)
    capi.loss(capi.LOSS_MULTI, fp.flat, B, cc, ss, 1, n_cmp, d, 1.0, beta, 1e-3, grad_out=g_32, prec=capi.PREC_FP32)
    P = {k: v.detach().cpu().clone() for k, v in fp.views.items()}
    _, _, _, grads = O.loss_and_grads(P, cc.cpu().unsqueeze(0), ss.cpu().unsqueeze(0), B.cpu(), beta, 1e-3, "multi")
    per = {}
    for k in P:
        if grads[k] is None: continue
        o, n = fp.offsets[k], P[k].numel()
        go = grads[k].reshape(-1).double(); a = g_tc[o:o + n].cpu().double(); b = g_32[o:o + n].cpu().double()
        per[k] = {"tc_max": float((a - go).abs().max() / go.abs().max()), "tc_l2": float(torch.linalg.norm(a - go) / torch.linalg.norm(go)),
                  "f32_max": float((b - go).abs().max() / go.abs().max())}
    # stats at the full training batch too (largest adjoints)
    g_tmp = torch.zeros_like(fp.flat)
    capi.loss(capi.LOSS_MULTI, fp.flat, B, c, s, 1, n_train, d, 1.0, beta, 1e-3, grad_out=g_tmp, prec=capi.PREC_TC)
    st_full = capi.adjoint_stats(1, n_train, d)
    print(json.dumps({"label": "trained_grads", "d": d, "loss_first": hist[0], "loss_last": hist[-1], "beta": beta,
                      "adjoint_stats_cmp": st, "adjoint_stats_full": st_full,
                      "worst_tc_max": max(v["tc_max"] for v in per.values()), "worst_tc_l2": max(v["tc_l2"] for v in per.values()),
                      "worst_f32_max": max(v["f32_max"] for v in per.values()), "per": per}), flush=True)

if __name__ == "__main__":
    forward_stats(3, 1, 100000, 0.5 / 5, "tc", "configs[0] test box E=1")
    forward_stats(3, 2, 50000, 0.5 / 5, "tc", "configs[0] test box E=2")
    forward_stats(6, 1, 20000, 0.5 / 12, "tc", "configs[3] UR5 slice")
    forward_stats(3, 1, 20000, 0.5 / 5, "fp32", "fp32 verify d3", arb64=True)
    forward_stats(6, 1, 5000, 0.5 / 12, "fp32", "fp32 verify d6", arb64=True)
    trained_grads(3)
    trained_grads(6, steps=200, n_train=12000, n_cmp=200)
