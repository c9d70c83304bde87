"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol the
header declares, the parameter layout matches the reference's state_dict, argument errors are
reported without touching a GPU, and the data-parallel host logic works under gloo (world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from pntf import capi
from pntf.params import TENSOR_NAMES, tensor_offsets, tensor_shapes
from oracle import pntf_oracle as O


def header_functions():
    src = open(os.path.join(ROOT, "include", "pntf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pntf_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = capi.lib()
    names = header_functions()
    assert len(names) >= 11
    for n in names:
        assert hasattr(L, n), n
    assert sorted(capi.EXPORTS) == names
    assert L.pntf_version() >= 100


def test_param_layout_matches_state_dict_order():
    table = capi.param_table()
    offs, shapes = tensor_offsets(), tensor_shapes()
    sd = O.init_params(0)
    assert list(sd.keys()) == list(TENSOR_NAMES)
    total = 0
    for i, k in enumerate(TENSOR_NAMES):
        off, rows, cols, used = table[i]
        assert off == offs[k] == total
        assert rows * cols == int(np.prod(shapes[k])) == sd[k].numel()
        assert tuple(sd[k].shape) == tuple(shapes[k])
        assert used == (0 if k.startswith("encoder1.0.") else 1)
        total += sd[k].numel()
    assert total == capi.PARAM_COUNT == 576129


def test_argument_errors_without_gpu():
    L = capi.lib()
    dummy = ctypes.c_void_p(0x1000)
    rc = L.pntf_forward(capi.MODE_LAPLACE, 0, dummy, dummy, dummy, 1, 8, 5, dummy, dummy, dummy, 0, dummy, 1 << 30, None)
    assert rc == 2 and b"d=5" in L.pntf_last_error()
    rc = L.pntf_forward(7, 0, dummy, dummy, dummy, 1, 8, 3, dummy, dummy, dummy, 0, dummy, 1 << 30, None)
    assert rc == 1
    rc = L.pntf_forward(capi.MODE_GRAD, 0, None, dummy, dummy, 1, 8, 3, dummy, dummy, dummy, 0, dummy, 1 << 30, None)
    assert rc == 1
    rc = L.pntf_loss_fwd(3, dummy, dummy, dummy, dummy, 1, 8, 3, 1.0, 1.0, 1e-3, dummy, dummy, 0, dummy, 1 << 30, None)
    assert rc == 1
    # the tensor-core mode validates before dispatching as well (d outside {3, 6}, null pointers, E < 1, bad prec)
    TC = capi.PREC_TC
    rc = L.pntf_forward(capi.MODE_GRAD, 0, dummy, dummy, dummy, 1, 8, 4, dummy, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 2 and b"d=4" in L.pntf_last_error()
    rc = L.pntf_forward(capi.MODE_GRAD, 0, dummy, None, dummy, 1, 8, 3, dummy, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 1
    rc = L.pntf_forward(capi.MODE_GRAD, 0, dummy, dummy, dummy, 0, 8, 3, dummy, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 1
    rc = L.pntf_loss_fwd(1, dummy, dummy, dummy, dummy, 1, 8, 5, 1.0, 1.0, 1e-3, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 2
    rc = L.pntf_loss_fwd(1, None, dummy, dummy, dummy, 1, 8, 3, 1.0, 1.0, 1e-3, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 1
    rc = L.pntf_loss_bwd(1, dummy, dummy, dummy, dummy, 1, 8, 7, 1.0, 1.0, 1e-3, dummy, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 2
    rc = L.pntf_field(capi.FIELD_SPEED, 0, dummy, dummy, dummy, 1, 8, 2, 0.0, dummy, TC, dummy, 1 << 30, None)
    assert rc == 2
    rc = L.pntf_plan(dummy, dummy, dummy, 8, 9, 0.03, 0.06, 10, 0, dummy, dummy, TC, dummy, 1 << 30, None)
    assert rc == 2
    rc = L.pntf_forward(capi.MODE_GRAD, 0, dummy, dummy, dummy, 1, 8, 3, dummy, dummy, dummy, 5, dummy, 1 << 30, None)
    assert rc == 1 and b"precision" in L.pntf_last_error()
    assert capi.lib().pntf_workspace_bytes(capi.OP_FORWARD, 1, 100, 4, 0, 0) == 0
    with pytest.raises(RuntimeError):
        capi.forward(capi.MODE_VALUE, torch.zeros(capi.PARAM_COUNT), torch.zeros(3, 128), torch.zeros(4, 6), 1, 4, 3)


def test_workspace_sizes():
    small = capi.workspace_bytes(capi.OP_FORWARD, 1, 1000, 3)
    assert 2_000_000 < small < 3_000_000
    full = capi.workspace_bytes(capi.OP_LOSS_BWD, 1, 1 << 20, 3)
    assert full <= (8 << 30) + (8 << 20)
    tiny = capi.workspace_bytes(capi.OP_LOSS_BWD, 1, 64, 3)
    assert tiny < 64 * 1024 * 1024
    assert capi.workspace_bytes(capi.OP_LOSS_BWD, 1, 64, 6) > tiny


def test_drop_in_modules_and_state_dict_keys():
    from models import model_res_sigmoid as S, model_res_sigmoid_multi as M
    net = M.NN("cpu", 3)
    assert list(net.state_dict().keys()) == list(TENSOR_NAMES)
    net.load_state_dict(O.init_params(1), strict=True)
    net.apply(net.init_weights)
    w = net.encoder[0].weight
    assert float(w.abs().max()) <= 2.0 / np.sqrt(256) + 1e-6
    nets = S.NN("cpu", 6, torch.zeros(128, 6))
    assert tuple(nets.B.shape) == (6, 128)
    with pytest.raises(RuntimeError):          # no CPU fallback
        net.out(torch.zeros(4, 6), torch.zeros(3, 128))
    m = M.Model("/tmp", "/tmp", 3, 2, device="cpu")
    assert m.Params['Training']['Batch Size'] == 2 and m.len == 2
    s = S.Model("/tmp", "/tmp", 6, device="cpu")
    assert s.Params['Training']['Batch Size'] == 10000


def test_shard_range_partitions():
    from pntf.dist import shard_range
    for n in (0, 1, 7, 10000, 1 << 20):
        for ws in (1, 2, 3, 8):
            edges = [shard_range(n, r, ws) for r in range(ws)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            for a, b in zip(edges[:-1], edges[1:]):
                assert a[1] == b[0]
            sizes = [e - b for b, e in edges]
            assert max(sizes) - min(sizes) <= 1


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.join(%(root)r, "p-ntfields_b200"))
from pntf.dist import GradAllReduce, shard_range, world
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
rank, ws = world()
red = GradAllReduce(1000, torch.device("cpu"))
g = torch.full((1000,), float(rank + 1))
sc = torch.tensor([float(rank), 10.0 * (rank + 1)])
g, sc = red(g, sc)
assert torch.allclose(g, torch.full((1000,), 3.0)), g[:4]
assert torch.allclose(sc, torch.tensor([0.5, 15.0])), sc
b, e = shard_range(10001, rank, ws)
t = torch.tensor([float(e - b)]); dist.all_reduce(t)
assert t.item() == 10001
dist.destroy_process_group()
print("ok", rank)
"""


def test_gradient_allreduce_gloo_world2(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER % {"root": ROOT, "port": port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                              text=True) for r in range(2)]
    outs = [p.communicate(timeout=120)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


def test_bench_reads_committed_traffic_table_and_parses_power_samples(tmp_path):
    """bench.py's roofline.traffic comes from the newest committed ncu capture (no constants in bench.py), and the clocks
    block is computed from nvidia-smi samples: under-load median clock, throttle reasons, averaged and instantaneous board
    power."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("pntf_bench", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    table, src = bench.dram_bytes_per_tile()
    assert src is not None and src.startswith("profiles/") and src.endswith("dram_bytes_per_tile.json")
    for k in ("tc_forward", "tc_reverse", "tc_dw"):
        assert 2e5 < table[k] < 2e6, (k, table[k])          # a few hundred KB per 128-row tile and kernel
    s = bench.ClockSampler(0)
    s.path = str(tmp_path / "clocks.csv")
    with open(s.path, "w") as f:
        f.write("0, 1965, 1965, 180.0, 0x0, Not Active, Not Active, Not Active, Not Active, 185.0\n")
        f.write("0, 1620, 1965, 590.0, 0x4, Not Active, Not Active, Not Active, Active, 980.0\n")
        f.write("0, 1590, 1965, 600.0, 0x4, Not Active, Not Active, Not Active, Active, [N/A]\n")   # no instantaneous reading in this sample
        f.write("garbage line\n")

    class _P:
        def terminate(self): pass
        def wait(self, timeout=None): return 0
        def kill(self): pass
    s.proc, s.f = _P(), open(os.devnull, "w")
    c = s.stop()
    assert c["sm_mhz"] == 1605.0 and c["sm_max_mhz"] == 1965.0 and c["reasons"] == ["sw_power_cap"]
    assert c["power_w_max"] == 600.0 and c["power_w_instant_max"] == 980.0 and c["samples"] == 3
