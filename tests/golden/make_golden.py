"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

Run (build container only; /root/reference does not exist on the GPU box):
    python tests/golden/make_golden.py

The reference modules are imported as they are through a 2-module shim (empty ``matplotlib`` /
``matplotlib.pylab`` and ``pickle5`` -> stdlib ``pickle``).  Weights come
from ``oracle.pntf_oracle.init_params(seed)`` (numpy RandomState, reproducible everywhere) and are
loaded into the reference ``NN`` with ``load_state_dict(strict=True)``; inputs are stored in the
fixture.  Every reference output that the tests compare against is stored as float32.
"""
import os
import pickle
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

for _m in ("matplotlib", "matplotlib.pylab"):
    sys.modules.setdefault(_m, types.ModuleType(_m))
sys.modules.setdefault("pickle5", pickle)
sys.path.insert(0, "/root/reference")
from models import model_res_sigmoid as S          # noqa: E402
from models import model_res_sigmoid_multi as M    # noqa: E402

from oracle import pntf_oracle as O                # noqa: E402

torch.set_num_threads(8)


def make_pairs(rs, n, d, lo=-0.5, hi=0.5):
    """Start/goal pairs in the unit box, rounded through fp16 like models/data_multi.py:19."""
    x = rs.uniform(lo, hi, size=(n, 2 * d)).astype(np.float32)
    return x.astype(np.float16).astype(np.float32)


def make_speed(rs, n):
    return rs.uniform(0.1, 1.0, size=(n, 2)).astype(np.float32)


def digest(name, t):
    """Compact fixture of a parameter-shaped tensor: biases whole; weights as 4 full rows,
    row sums, column sums and the Frobenius norm (keeps the committed fixtures small)."""
    t = np.asarray(t, np.float32)
    if t.ndim == 1:
        return {name: t}
    fo = t.shape[0]
    rows = sorted(set([0, min(1, fo - 1), fo // 2, fo - 1]))
    return {name + "#rows": t[rows], name + "#rowidx": np.asarray(rows, np.int32),
            name + "#rowsum": t.sum(1), name + "#colsum": t.sum(0),
            name + "#fro": np.float32(np.sqrt((t.astype(np.float64) ** 2).sum()))}


def grads_of(net):
    out = {}
    for k, p in net.named_parameters():
        out[k] = None if p.grad is None else p.grad.detach().clone()
    return out


def run_multi(seed, d, E, N):
    rs = np.random.RandomState(1000 + seed)
    P = O.init_params(seed)
    B = (0.5 * rs.normal(size=(E, d, 128))).astype(np.float32)
    coords = np.stack([make_pairs(rs, N, d) for _ in range(E)])
    speed = np.stack([make_speed(rs, N) for _ in range(E)])
    beta, gamma = 1.7, 1e-3

    model = M.Model("/tmp", "/tmp", d, E, device="cpu")
    net = M.NN("cpu", d)
    net.load_state_dict(P, strict=True)
    model.network = net
    tB, tc, ts = torch.from_numpy(B), torch.from_numpy(coords), torch.from_numpy(speed)
    fx = dict(seed=seed, d=d, E=E, N=N, B=B, coords=coords, speed=speed, beta=beta, gamma=gamma)

    with torch.no_grad():
        tau0, _ = net.out(tc[0], tB[0])
        tau1, dtau1, _ = net.out_grad(tc[0], tB[0])
        tau2, dtau2, ltau2, _ = net.out_laplace(tc, tB)
        _, dback, _ = net.out_backgrad(tc[0], tB[0])
    fx.update(out_tau=tau0.numpy(), grad_tau=tau1.numpy(), grad_dtau=dtau1.numpy(),
              lap_tau=tau2.numpy(), lap_dtau=dtau2.numpy(), lap_ltau=ltau2.numpy(),
              backgrad_dtau=dback.numpy())

    loss, loss_n, diff = model.Loss(tc, ts, tB, beta, gamma)
    loss.backward()
    g = grads_of(net)
    fx.update(loss=np.float32(loss.item()), loss_n=np.float32(loss_n.item()), diff=diff.detach().numpy())
    assert g["encoder1.0.weight"] is None and g["encoder1.0.bias"] is None
    for k, v in g.items():
        if v is not None:
            fx.update(digest("grad/" + k, v.numpy()))

    # one AdamW step exactly as the trainer configures it (M:960-962)
    opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=0.1)
    opt.step()
    for k, v in net.state_dict().items():
        fx.update(digest("adamw/" + k, v.detach().numpy().copy()))
    net.load_state_dict(P, strict=True)

    with torch.no_grad():
        model.B = tB[0]
        fx["travel_times"] = model.TravelTimes(tc[0]).numpy()
        fx["speed_pred"] = model.Speed(tc[0]).numpy()
        fx["gradient"] = model.Gradient(tc[0], tB[0]).numpy()

    # plan loop, one query at a time as test/gib_plan.py:53-86 does (thr 0.06, step 0.03, cap 500)
    Q = 6
    sg = rs.uniform(-0.4, 0.4, size=(Q, 2 * d)).astype(np.float32)
    cap = 500
    ways = np.zeros((Q, cap + 2, 2 * d), np.float32)
    nst = np.zeros(Q, np.int32)
    with torch.no_grad():
        for q in range(Q):
            XP = torch.from_numpy(sg[q:q + 1].copy())
            dis = torch.norm(XP[:, d:] - XP[:, :d])
            ways[q, :, :] = XP.numpy()
            it = 0
            while dis > 0.06:
                gradient = model.Gradient(XP.clone(), tB[0])
                XP = XP + 0.03 * gradient
                dis = torch.norm(XP[:, d:] - XP[:, :d])
                it += 1
                ways[q, it:, :] = XP.numpy()
                if it > cap:
                    break
            nst[q] = it
    fx.update(plan_start_goal=sg, plan_waypoints=ways[:, :int(nst.max()) + 1].copy(), plan_nsteps=nst,
              plan_step=0.03, plan_thr=0.06, plan_cap=cap)
    return fx


def run_single(seed, d, N):
    rs = np.random.RandomState(2000 + seed)
    P = O.init_params(seed)
    Bsingle = (0.5 * rs.normal(size=(128, d))).astype(np.float32)     # S:942 layout [128,d]
    coords = make_pairs(rs, N, d)
    speed = make_speed(rs, N)
    beta, gamma = 0.9, 1e-3

    model = S.Model("/tmp", "/tmp", d, device="cpu")
    net = S.NN("cpu", d, torch.from_numpy(Bsingle))
    net.load_state_dict(P, strict=True)
    model.network = net
    tc, ts = torch.from_numpy(coords), torch.from_numpy(speed)
    fx = dict(seed=seed, d=d, N=N, B=Bsingle, coords=coords, speed=speed, beta=beta, gamma=gamma)

    tau0, _ = net.out(tc)
    with torch.no_grad():
        tau1, dtau1, _ = net.out_grad(tc)
        tau2, dtau2, ltau2, _ = net.out_laplace(tc)
    fx.update(out_tau=tau0.detach().numpy(), grad_tau=tau1.numpy(), grad_dtau=dtau1.numpy(),
              lap_tau=tau2.numpy(), lap_dtau=dtau2.numpy(), lap_ltau=ltau2.numpy())

    loss, loss_n, diff = model.Loss(tc, ts, beta, gamma)
    loss.backward()
    g = grads_of(net)
    fx.update(loss=np.float32(loss.item()), loss_n=np.float32(loss_n.item()), diff=diff.detach().numpy())
    for k, v in g.items():
        if v is not None:
            fx.update(digest("grad/" + k, v.numpy()))
    net.zero_grad()

    with torch.no_grad():
        fx["travel_times"] = model.TravelTimes(tc).numpy()
        fx["speed_pred"] = model.Speed(tc).numpy()
        fx["speed2_pred"] = model.Speed2(tc, gamma).numpy()
    # Model.Gradient of the single model is batch-1 only (S:1264-1280): call it row by row.
    rows = []
    for i in range(min(N, 8)):
        rows.append(model.Gradient(tc[i:i + 1].clone()).detach().numpy()[0])
    fx["gradient_rows"] = np.stack(rows)

    # plan loop as test/arm_plan.py:111-152 (thr 0.03, step 0.015, cap 300)
    Q = 3
    sg = rs.uniform(-0.4, 0.4, size=(Q, 2 * d)).astype(np.float32)
    cap = 300
    ways = np.zeros((Q, cap + 2, 2 * d), np.float32)
    nst = np.zeros(Q, np.int32)
    for q in range(Q):
        XP = torch.from_numpy(sg[q:q + 1].copy())
        dis = torch.norm(XP[:, d:] - XP[:, :d])
        ways[q, :, :] = XP.numpy()
        it = 0
        while dis > 0.03:
            gradient = model.Gradient(XP.clone()).detach()
            XP = XP + 0.015 * gradient
            dis = torch.norm(XP[:, d:] - XP[:, :d])
            it += 1
            ways[q, it:, :] = XP.numpy()
            if it > cap:
                break
        nst[q] = it
    fx.update(plan_start_goal=sg, plan_waypoints=ways[:, :int(nst.max()) + 1].copy(), plan_nsteps=nst,
              plan_step=0.015, plan_thr=0.03, plan_cap=cap)
    return fx


def run_adamw_kat():
    """Known-answer vectors for the optimiser: torch.optim.AdamW(lr=1e-3, weight_decay=0.1)
    (S:954-956) on one flat parameter, three steps with stored gradients."""
    rs = np.random.RandomState(77)
    n = 4096
    p0 = rs.uniform(-0.2, 0.2, size=n).astype(np.float32)
    gs = (rs.normal(size=(3, n)) * np.array([1e-2, 1e-5, 1.0])[:, None]).astype(np.float32)
    p = torch.nn.Parameter(torch.from_numpy(p0.copy()))
    opt = torch.optim.AdamW([p], lr=1e-3, weight_decay=0.1)
    fx = dict(p0=p0, grads=gs, lrs=np.array([1e-3, 1e-3, 5e-4], np.float64))
    for t in range(3):
        opt.param_groups[0]["lr"] = float(fx["lrs"][t])
        p.grad = torch.from_numpy(gs[t].copy())
        opt.step()
        fx["p%d" % (t + 1)] = p.detach().numpy().copy()
    st = opt.state[p]
    fx["m3"] = st["exp_avg"].numpy().copy()
    fx["v3"] = st["exp_avg_sq"].numpy().copy()
    return fx


def save(name, fx):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in fx.items()})
    print(name, "%.1f KB" % (os.path.getsize(path) / 1e3), "keys", len(fx))


if __name__ == "__main__":
    save("adamw_kat.npz", run_adamw_kat())
    save("multi_d3.npz", run_multi(seed=11, d=3, E=2, N=40))
    save("single_d6.npz", run_single(seed=12, d=6, N=24))
    save("single_d3.npz", run_single(seed=13, d=3, N=24))
