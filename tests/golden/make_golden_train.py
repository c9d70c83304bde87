"""Golden fixtures of the host-side rows: trainer schedule (T2), checkpoint interop and field slices (8f rank 4).

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_train.py

The UNMODIFIED reference trainers (models/model_res_sigmoid_multi.py:953-1141, models/model_res_sigmoid.py:938-1137)
are run on CPU on a small synthetic dataset written in the reference's on-disk layout.  Nothing in the reference is
edited; the run is *observed* by wrapping library functions it calls:
  torch.randperm        -> every permutation it draws is recorded (DataLoader shuffle and row permutations)
  torch.optim.AdamW     -> at construction (right after init_weights) the parameters are overwritten with
                           oracle.init_params(seed), so the run starts from weights the tests can rebuild
  Model.Loss            -> beta, gamma, lr, loss, loss_n of every step
  random.randint        -> the rollback draws
  matplotlib stub       -> the arrays Model.plot hands to pcolormesh / contour (X, Y, V / TAU, TT)
Outputs: train_multi_d3.npz, train_single_d6.npz, plot_slices.npz, ref_ckpt/Model_Epoch_*.pt (reference-written).
"""
import contextlib
import io
import os
import pickle
import random
import shutil
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

PLOT_CALLS = []


class _Anything:
    """Swallows every attribute access / call (figure, axes, colorbar, savefig ...); records the array arguments of
    pcolormesh and contour."""

    def __init__(self, name="plt"):
        self._name = name

    def __getattr__(self, k):
        if k.startswith("__"):
            raise AttributeError(k)
        return _Anything(self._name + "." + k)

    def __call__(self, *a, **kw):
        leaf = self._name.rsplit(".", 1)[-1]
        if leaf in ("pcolormesh", "contour"):
            PLOT_CALLS.append((leaf, [np.asarray(x).copy() for x in a[:3]]))
        return _Anything(self._name + "()")


mpl = types.ModuleType("matplotlib")
pylab = types.ModuleType("matplotlib.pylab")
_plt = _Anything()
def _pylab_getattr(k):
    if k.startswith("__"):
        raise AttributeError(k)
    return getattr(_plt, k)


pylab.__getattr__ = _pylab_getattr
mpl.pylab = pylab
sys.modules["matplotlib"] = mpl
sys.modules["matplotlib.pylab"] = pylab
sys.modules.setdefault("pickle5", pickle)
sys.path.insert(0, "/root/reference")
from models import model_res_sigmoid as S          # noqa: E402
from models import model_res_sigmoid_multi as M    # noqa: E402

from oracle import pntf_oracle as O                # noqa: E402
import helpers                                     # noqa: E402

torch.set_num_threads(8)


class Recorder:
    def __init__(self, init_seed):
        self.perms, self.randints, self.steps = [], [], []
        self.init_seed = init_seed

    @contextlib.contextmanager
    def active(self, module):
        rec = self
        orig_randperm, orig_randint = torch.randperm, random.randint
        orig_init, orig_loss = torch.optim.AdamW.__init__, module.Model.Loss

        def randperm(n, *a, **kw):
            r = orig_randperm(n, *a, **kw)
            rec.perms.append(r.clone())
            return r

        def randint(a, b):
            v = orig_randint(a, b)
            rec.randints.append(v)
            return v

        def adamw_init(self_, params, *a, **kw):
            params = list(params)
            P = O.init_params(rec.init_seed)
            vals = list(P.values())
            assert len(vals) == len(params)
            with torch.no_grad():
                for p, v in zip(params, vals):
                    p.copy_(v)
            return orig_init(self_, params, *a, **kw)

        def loss(self_, *a):
            out = orig_loss(self_, *a)
            beta, gamma = a[-2], a[-1]
            rec.steps.append((float(beta), float(gamma), float(self_.optimizer.param_groups[0]["lr"]),
                              float(out[0].item()), float(out[1].item()), float(a[1].reshape(-1)[0].item())))
            return out

        torch.randperm, random.randint = randperm, randint
        torch.optim.AdamW.__init__, module.Model.Loss = adamw_init, loss
        try:
            yield
        finally:
            torch.randperm, random.randint = orig_randperm, orig_randint
            torch.optim.AdamW.__init__, module.Model.Loss = orig_init, orig_loss


def digest_state(sd):
    out = {}
    for k, v in sd.items():
        t = v.detach().numpy().astype(np.float32)
        out["final/" + k + "#fro"] = np.float32(np.sqrt((t.astype(np.float64) ** 2).sum()))
        out["final/" + k + "#head"] = t.reshape(-1)[:16].copy()
    return out


def parse_log(text):
    ep = []
    for line in text.splitlines():
        if line.startswith("Epoch = "):
            parts = line.replace("--", "").split()
            ep.append((int(parts[2]), float(parts[5]), float(parts[8])))
    return ep


def run_multi_train(seed=21, d=3, rows=12000, epochs=8):
    tmp = tempfile.mkdtemp(prefix="pntf_golden_")
    data_root = os.path.join(tmp, "data") + "/"
    model_path = os.path.join(tmp, "model")
    os.makedirs(model_path)
    envs = [helpers.train_fixture_env(seed + i, rows, d) for i in range(2)]
    helpers.write_multi_dataset(data_root, envs)
    model = M.Model(model_path, data_root, d, 2, device="cpu")
    model.Params["Training"]["Number of Epochs"] = epochs
    rec = Recorder(seed)
    torch.manual_seed(seed)
    random.seed(seed)
    buf = io.StringIO()
    with rec.active(M), contextlib.redirect_stdout(buf):
        model.train()
    log = buf.getvalue()
    print(log)
    # draws per while-iteration: DataLoader order (n=2), idx0, idx1 (n=rows)
    assert all(p.numel() in (2, rows) for p in rec.perms)
    # torch's RandomSampler draws randperm(n) twice per pass: the order it yields, then one more for the
    # (empty) remainder `[:num_samples % n]`; the environment order of a pass is the first of each pair
    env_orders = np.stack([p.numpy() for p in rec.perms if p.numel() == 2]).astype(np.uint8)[::2]
    row_perms = np.stack([p.numpy()[:10000] for p in rec.perms if p.numel() == rows]).astype(np.uint16)
    steps = np.asarray(rec.steps, np.float64)
    fx = dict(seed=seed, d=d, rows=rows, epochs=epochs, env_orders=env_orders, row_perms=row_perms, steps=steps,
              epoch_log=np.asarray(parse_log(log), np.float64), randints=np.asarray(rec.randints, np.int32),
              total_train_loss=np.asarray([float(x) for x in model.total_train_loss], np.float64),
              files=np.asarray(sorted(os.listdir(model_path))), final_B=model.B.numpy())
    fx.update(digest_state(model.network.state_dict()))
    shutil.rmtree(tmp)
    return fx


def run_single_train(seed=31, d=6, rows=12000, epochs=3):
    tmp = tempfile.mkdtemp(prefix="pntf_golden_")
    data_root = os.path.join(tmp, "data")
    model_path = os.path.join(tmp, "model")
    os.makedirs(model_path)
    pts, spd, _ = helpers.train_fixture_env(seed, rows, d)
    helpers.write_single_dataset(data_root, pts, spd)
    model = S.Model(model_path, data_root, d, device="cpu")
    model.Params["Training"]["Number of Epochs"] = epochs
    rec = Recorder(seed)
    torch.manual_seed(seed)
    random.seed(seed)
    buf = io.StringIO()
    with rec.active(S), contextlib.redirect_stdout(buf):
        model.train()
    log = buf.getvalue()
    print(log)
    assert all(p.numel() == rows for p in rec.perms)
    fx = dict(seed=seed, d=d, rows=rows, epochs=epochs, row_perms=np.stack([p.numpy() for p in rec.perms]).astype(np.uint16),
              steps=np.asarray(rec.steps, np.float64), epoch_log=np.asarray(parse_log(log), np.float64),
              randints=np.asarray(rec.randints, np.int32), B=model.B.numpy(),
              total_train_loss=np.asarray([float(x) for x in model.total_train_loss], np.float64),
              files=np.asarray(sorted(os.listdir(model_path))))
    fx.update(digest_state(model.network.state_dict()))
    shutil.rmtree(tmp)
    return fx


def run_plot_slices(seed=41):
    """Model.plot (M:1250-1293, S:1284-1329) on weights from oracle.init_params: the arrays it draws."""
    fx = {}
    P = O.init_params(seed)
    rs = np.random.RandomState(seed)
    # multi, d = 3
    Bm = (0.5 * rs.normal(size=(3, 128))).astype(np.float32)
    model = M.Model("/tmp", "/tmp", 3, 2, device="cpu")
    model.network = M.NN("cpu", 3)
    model.network.load_state_dict(P, strict=True)
    model.B = torch.from_numpy(Bm)
    del PLOT_CALLS[:]
    with torch.no_grad():
        model.plot(1, 0.5, 0.5)
    calls = list(PLOT_CALLS)
    assert [c[0] for c in calls] == ["pcolormesh", "contour", "pcolormesh", "contour"]
    fx.update(seed=seed, multi_B=Bm, multi_X=calls[0][1][0], multi_Y=calls[0][1][1], multi_V=calls[0][1][2],
              multi_TT=calls[1][1][2], multi_TAU=calls[2][1][2])
    # single, d = 6
    Bs = (0.5 * rs.normal(size=(128, 6))).astype(np.float32)
    model = S.Model("/tmp", "/tmp", 6, device="cpu")
    model.network = S.NN("cpu", 6, torch.from_numpy(Bs))
    model.network.load_state_dict(P, strict=True)
    del PLOT_CALLS[:]
    model.plot(1, 0.5, 0.5)
    calls = list(PLOT_CALLS)
    fx.update(single_B=Bs, single_X=calls[0][1][0], single_Y=calls[0][1][1], single_V=calls[0][1][2],
              single_TT=calls[1][1][2], single_TAU=calls[2][1][2])
    return fx


def write_reference_checkpoint(seed=51):
    """A checkpoint written by the reference's own Model.save (M:1143-1152) from known weights (optimiser constructed,
    no step taken, so its state is empty and the file stays at the size of the weights)."""
    out = os.path.join(HERE, "ref_ckpt")
    shutil.rmtree(out, ignore_errors=True)
    os.makedirs(out)
    model = M.Model(out, "/tmp", 3, 2, device="cpu")
    model.network = M.NN("cpu", 3)
    model.network.load_state_dict(O.init_params(seed), strict=True)
    model.optimizer = torch.optim.AdamW(model.network.parameters(), lr=1e-3, weight_decay=0.1)
    rs = np.random.RandomState(seed)
    model.B = torch.from_numpy((0.5 * rs.normal(size=(3, 128))).astype(np.float32))
    model.total_train_loss = [torch.tensor(0.25), torch.tensor(0.125)]
    model.save(epoch=3, val_loss=torch.tensor(0.0125))
    print("ref_ckpt:", os.listdir(out))


def save(name, fx):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in fx.items()})
    print(name, "%.1f KB" % (os.path.getsize(path) / 1e3), "keys", len(fx))


if __name__ == "__main__":
    what = sys.argv[1:] or ["plot", "ckpt", "multi", "single"]
    if "plot" in what:
        save("plot_slices.npz", run_plot_slices())
    if "ckpt" in what:
        write_reference_checkpoint()
    if "multi" in what:
        save("train_multi_d3.npz", run_multi_train())
    if "single" in what:
        save("train_single_d6.npz", run_single_train())
