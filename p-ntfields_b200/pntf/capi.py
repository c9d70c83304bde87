"""ctypes binding of libpntf_b200.so (C-ABI: include/pntf_b200.h).

Every wrapper takes torch CUDA tensors, checks dtype/contiguity/device, passes raw device pointers
and the current torch CUDA stream, and raises RuntimeError on a nonzero status.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpntf_b200.so")

PARAM_COUNT = 576129
PREC_FP32, PREC_TC = 0, 1
MODE_VALUE, MODE_GRAD, MODE_LAPLACE = 0, 1, 2
LOSS_SINGLE, LOSS_MULTI = 0, 1
FIELD_TRAVEL_TIME, FIELD_SPEED, FIELD_SPEED2, FIELD_DIRECTION = 0, 1, 2, 3
GRAD_EXACT, GRAD_COMPAT = 0, 1
OP_FORWARD, OP_LOSS_FWD, OP_LOSS_BWD, OP_PLAN, OP_FIELD = 0, 1, 2, 3, 4

EXPORTS = ["pntf_version", "pntf_last_error", "pntf_param_offset", "pntf_workspace_bytes", "pntf_forward",
           "pntf_field", "pntf_loss_fwd", "pntf_loss_bwd", "pntf_loss_fwd_gather", "pntf_loss_bwd_gather", "pntf_loss_bwd_gather_sched", "pntf_adamw", "pntf_adamw_sched", "pntf_plan", "pntf_plan_compact", "pntf_plan_static", "pntf_launch_count", "pntf_profile_begin", "pntf_profile_end",
           "pntf_tc_selftest", "pntf_mesh_workspace_bytes", "pntf_mesh_distance", "pntf_debug_adjoint_stats", "pntf_set_deterministic"]

_lib = None


def lib():
    """Load the shared library (built by p-ntfields_b200/build.py / __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libpntf_b200.so is missing (%s): run `python p-ntfields_b200/build.py`; "
                           "there is no CPU fallback" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    c_i, c_i64, c_f, c_p, c_sz = ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_void_p, ctypes.c_size_t
    L.pntf_version.restype = c_i
    L.pntf_last_error.restype = ctypes.c_char_p
    L.pntf_launch_count.restype = c_i64
    L.pntf_param_offset.argtypes = [c_i, ctypes.POINTER(c_i64), ctypes.POINTER(c_i64), ctypes.POINTER(c_i64),
                                    ctypes.POINTER(c_i)]
    L.pntf_workspace_bytes.restype = c_sz
    L.pntf_workspace_bytes.argtypes = [c_i, c_i, c_i64, c_i, c_i, c_sz]
    L.pntf_forward.argtypes = [c_i, c_i, c_p, c_p, c_p, c_i, c_i64, c_i, c_p, c_p, c_p, c_i, c_p, c_sz, c_p]
    L.pntf_field.argtypes = [c_i, c_i, c_p, c_p, c_p, c_i, c_i64, c_i, c_f, c_p, c_i, c_p, c_sz, c_p]
    L.pntf_loss_fwd.argtypes = [c_i, c_p, c_p, c_p, c_p, c_i, c_i64, c_i, c_f, c_f, c_f, c_p, c_p, c_i, c_p, c_sz, c_p]
    L.pntf_loss_bwd.argtypes = [c_i, c_p, c_p, c_p, c_p, c_i, c_i64, c_i, c_f, c_f, c_f, c_p, c_p, c_p, c_i, c_p,
                                c_sz, c_p]
    L.pntf_loss_fwd_gather.argtypes = [c_i, c_p, c_p, c_p, c_i64, c_p, c_i, c_i64, c_i, c_f, c_f, c_f, c_p, c_p, c_i, c_p,
                                       c_sz, c_p]
    L.pntf_loss_bwd_gather.argtypes = [c_i, c_p, c_p, c_p, c_i64, c_p, c_i, c_i64, c_i, c_f, c_f, c_f, c_p, c_p, c_p, c_i,
                                       c_p, c_sz, c_p]
    L.pntf_loss_bwd_gather_sched.argtypes = [c_i, c_p, c_p, c_p, c_i64, c_p, c_i, c_i64, c_i, c_p, c_p, c_p, c_p, c_i, c_p,
                                             c_sz, c_p]
    c_d = ctypes.c_double
    L.pntf_adamw.argtypes = [c_p, c_p, c_p, c_p, c_p, c_i64, c_d, c_d, c_d, c_d, c_d, c_i64, c_d, c_p]
    L.pntf_adamw_sched.argtypes = [c_p, c_p, c_p, c_p, c_p, c_i64, c_p, c_d, c_d, c_d, c_d, c_d, c_p]
    L.pntf_plan.argtypes = [c_p, c_p, c_p, c_i64, c_i, c_f, c_f, c_i, c_i, c_p, c_p, c_i, c_p, c_sz, c_p]
    L.pntf_plan_compact.argtypes = L.pntf_plan.argtypes
    L.pntf_plan_static.argtypes = [c_p, c_p, c_p, c_i64, c_i, c_f, c_f, c_i, c_i, c_p, c_p, c_p]
    L.pntf_tc_selftest.argtypes = [c_p, c_p, c_p, c_p, c_p]
    L.pntf_debug_adjoint_stats.argtypes = [c_p, c_sz, c_i, c_i64, c_i, c_i, c_p, c_p]
    L.pntf_mesh_workspace_bytes.restype = c_sz
    L.pntf_mesh_workspace_bytes.argtypes = [c_i64]
    L.pntf_mesh_distance.argtypes = [c_p, c_i64, c_p, c_i64, c_p, c_f, c_f, c_p, c_p, c_sz, c_p]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise RuntimeError("libpntf_b200: error %d: %s" % (rc, lib().pntf_last_error().decode()))


def _ptr(t, dtype=torch.float32, name="tensor"):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (no CPU fallback)" % name)
    if t.dtype != dtype or not t.is_contiguous():
        raise RuntimeError("%s must be contiguous %s" % (name, dtype))
    return t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


_ws_cache = {}


def workspace(nbytes, device):
    """A cached scratch buffer of at least nbytes on `device` (grown geometrically)."""
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = None
        _ws_cache.pop(key, None)
        buf = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def workspace_bytes(op, E, N, d, prec=PREC_FP32, budget=0):
    n = lib().pntf_workspace_bytes(op, E, N, d, prec, budget)
    if n == 0:
        raise RuntimeError("libpntf_b200: unsupported configuration (d=%d, prec=%d)" % (d, prec))
    return n


def launch_count():
    return int(lib().pntf_launch_count())


KERNEL_CLASSES = ("pack", "forward", "reverse", "dw", "finalize", "adamw", "plan", "tc_forward", "tc_reverse", "tc_dw", "tc_backward")


def profile_begin():
    _check(lib().pntf_profile_begin())


def profile_end():
    """{class: (ms, launches)} of everything launched since profile_begin()."""
    n = len(KERNEL_CLASSES)
    ms = (ctypes.c_double * n)()
    cnt = (ctypes.c_int64 * n)()
    _check(lib().pntf_profile_end(ms, cnt, n))
    return {KERNEL_CLASSES[i]: (ms[i], cnt[i]) for i in range(n)}


def param_table():
    """[(offset, rows, cols, used)] for the 30 state_dict tensors."""
    out = []
    o, r, c, u = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int()
    for i in range(30):
        _check(lib().pntf_param_offset(i, ctypes.byref(o), ctypes.byref(r), ctypes.byref(c), ctypes.byref(u)))
        out.append((o.value, r.value, c.value, u.value))
    return out


def forward(mode, params, B, coords, E, N, d, prec=PREC_FP32, grad_flavour=GRAD_EXACT):
    """tau[E*N], dtau[E*N,2d] or None, ltau[E*N,2d] or None."""
    dev = coords.device
    n = E * N
    tau = torch.empty(n, dtype=torch.float32, device=dev)
    dtau = torch.empty(n, 2 * d, dtype=torch.float32, device=dev) if mode >= MODE_GRAD else None
    ltau = torch.empty(n, 2 * d, dtype=torch.float32, device=dev) if mode >= MODE_LAPLACE else None
    ws = workspace(workspace_bytes(OP_FORWARD, E, N, d, prec), dev)
    _check(lib().pntf_forward(mode, grad_flavour, _ptr(params, name="params"), _ptr(B, name="B"),
                              _ptr(coords, name="coords"), E, N, d, _ptr(tau), _ptr(dtau), _ptr(ltau), prec,
                              ws.data_ptr(), ws.numel(), _stream()))
    return tau, dtau, ltau


def field(kind, params, B, coords, E, N, d, gamma=0.0, prec=PREC_FP32, grad_flavour=GRAD_EXACT):
    dev = coords.device
    n = E * N
    shape = (n, 2 * d) if kind == FIELD_DIRECTION else (n,)
    out = torch.empty(shape, dtype=torch.float32, device=dev)
    ws = workspace(workspace_bytes(OP_FIELD, E, N, d, prec), dev)
    _check(lib().pntf_field(kind, grad_flavour, _ptr(params, name="params"), _ptr(B, name="B"),
                            _ptr(coords, name="coords"), E, N, d, float(gamma), _ptr(out), prec, ws.data_ptr(),
                            ws.numel(), _stream()))
    return out


def loss(variant, params, B, coords, speed, E, N, d, alpha, beta, gamma, grad_out=None, prec=PREC_FP32,
         budget=0, diff=None, loss_out=None):
    """Returns (loss_out[2] device tensor, diff[E*N]); accumulates into grad_out when given."""
    dev = coords.device
    n = E * N
    if diff is None:
        diff = torch.empty(n, dtype=torch.float32, device=dev)
    if loss_out is None:
        loss_out = torch.empty(2, dtype=torch.float32, device=dev)
    common = (variant, _ptr(params, name="params"), _ptr(B, name="B"), _ptr(coords, name="coords"),
              _ptr(speed, name="speed"), E, N, d, float(alpha), float(beta), float(gamma), _ptr(diff), _ptr(loss_out))
    if grad_out is None:
        ws = workspace(workspace_bytes(OP_LOSS_FWD, E, N, d, prec), dev)
        _check(lib().pntf_loss_fwd(*common, prec, ws.data_ptr(), ws.numel(), _stream()))
    else:
        if grad_out.numel() != PARAM_COUNT:
            raise RuntimeError("grad_out must have %d elements" % PARAM_COUNT)
        ws = workspace(workspace_bytes(OP_LOSS_BWD, E, N, d, prec, budget), dev)
        _check(lib().pntf_loss_bwd(*common, _ptr(grad_out, name="grad_out"), prec, ws.data_ptr(), ws.numel(),
                                   _stream()))
    return loss_out, diff


def loss_gather(variant, params, B, dataset, index, E, N, d, alpha, beta, gamma, grad_out=None, prec=PREC_FP32,
                budget=0, diff=None, loss_out=None):
    """loss() on the batch `dataset[index]` without materialising it: dataset[rows, 2d+2] (coords ++ speed, the
    reference's table layout), index[E*N] int32 row numbers.  Returns (loss_out[2], diff[E*N])."""
    dev = dataset.device
    n = E * N
    if dataset.dim() != 2 or dataset.shape[1] != 2 * d + 2 or not dataset.is_contiguous():
        raise RuntimeError("dataset must be a contiguous [rows, 2d+2] tensor")
    if index.numel() != n:
        raise RuntimeError("index must have E*N = %d elements" % n)
    if diff is None:
        diff = torch.empty(n, dtype=torch.float32, device=dev)
    if loss_out is None:
        loss_out = torch.empty(2, dtype=torch.float32, device=dev)
    common = (variant, _ptr(params, name="params"), _ptr(B, name="B"), _ptr(dataset, name="dataset"),
              int(dataset.shape[0]), _ptr(index, torch.int32, "index"), E, N, d, float(alpha), float(beta), float(gamma),
              _ptr(diff), _ptr(loss_out))
    if grad_out is None:
        ws = workspace(workspace_bytes(OP_LOSS_FWD, E, N, d, prec), dev)
        _check(lib().pntf_loss_fwd_gather(*common, prec, ws.data_ptr(), ws.numel(), _stream()))
    else:
        if grad_out.numel() != PARAM_COUNT:
            raise RuntimeError("grad_out must have %d elements" % PARAM_COUNT)
        ws = workspace(workspace_bytes(OP_LOSS_BWD, E, N, d, prec, budget), dev)
        _check(lib().pntf_loss_bwd_gather(*common, _ptr(grad_out, name="grad_out"), prec, ws.data_ptr(), ws.numel(),
                                          _stream()))
    return loss_out, diff


def loss_gather_sched(variant, params, B, dataset, index, E, N, d, sched, grad_out, diff, loss_out, prec=PREC_TC, budget=0):
    """loss_gather(..., grad_out=...) with {alpha, beta, gamma} read from the device tensor `sched` (3 floats) at run
    time: capturable in a CUDA graph (no allocation: every buffer, incl. the cached workspace, must exist already)."""
    dev = dataset.device
    ws = workspace(workspace_bytes(OP_LOSS_BWD, E, N, d, prec, budget), dev)
    _check(lib().pntf_loss_bwd_gather_sched(variant, _ptr(params, name="params"), _ptr(B, name="B"),
                                            _ptr(dataset, name="dataset"), int(dataset.shape[0]),
                                            _ptr(index, torch.int32, "index"), E, N, d, _ptr(sched, name="sched"),
                                            _ptr(diff), _ptr(loss_out), _ptr(grad_out, name="grad_out"), prec,
                                            ws.data_ptr(), ws.numel(), _stream()))


def adamw_sched(params, grads, exp_avg, exp_avg_sq, mask, lr_step, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.1,
                grad_scale=1.0):
    """adamw() with {lr, step} read from the device tensor `lr_step` (2 floats) at run time."""
    _check(lib().pntf_adamw_sched(_ptr(params, name="params"), _ptr(grads, name="grads"), _ptr(exp_avg),
                                  _ptr(exp_avg_sq), _ptr(mask, torch.uint8, "mask"), params.numel(),
                                  _ptr(lr_step, name="lr_step"), float(betas[0]), float(betas[1]), float(eps),
                                  float(weight_decay), float(grad_scale), _stream()))


def adamw(params, grads, exp_avg, exp_avg_sq, mask, step, lr=1e-3, betas=(0.9, 0.999), eps=1e-8,
          weight_decay=0.1, grad_scale=1.0):
    n = params.numel()
    _check(lib().pntf_adamw(_ptr(params, name="params"), _ptr(grads, name="grads"), _ptr(exp_avg), _ptr(exp_avg_sq),
                            _ptr(mask, torch.uint8, "mask"), n, float(lr), float(betas[0]), float(betas[1]),
                            float(eps), float(weight_decay), int(step), float(grad_scale), _stream()))


def plan(params, B, start_goal, d, eta, thr, max_iter, grad_flavour=GRAD_EXACT, prec=PREC_FP32, fill_tail=True):
    """waypoints[Q, max_iter+2, 2d], nsteps[Q].  fill_tail=False (pntf_plan_compact): rows after nsteps[q] are not
    written (the buffer is torch.empty: read waypoints[q, :nsteps[q]+1] only)."""
    dev = start_goal.device
    Q = start_goal.shape[0]
    way = torch.empty(Q, max_iter + 2, 2 * d, dtype=torch.float32, device=dev)
    nsteps = torch.zeros(Q, dtype=torch.int32, device=dev)
    ws = workspace(workspace_bytes(OP_PLAN, 1, Q, d, prec), dev)
    fn = lib().pntf_plan if fill_tail else lib().pntf_plan_compact
    _check(fn(_ptr(params, name="params"), _ptr(B, name="B"), _ptr(start_goal, name="start_goal"), Q, d,
              float(eta), float(thr), int(max_iter), grad_flavour, _ptr(way),
              _ptr(nsteps, torch.int32), prec, ws.data_ptr(), ws.numel(), _stream()))
    return way, nsteps


PLAN_STATIC_MAX_QUERIES = 65536      # PNTF_PLAN_STATIC_MAX_QUERIES (include/pntf_b200.h)


def plan_static(params, B, start_goal, d, eta, thr, max_iter, way, nsteps, grad_flavour=GRAD_EXACT):
    """pntf_plan_static: the tensor-core planner with its scratch inside the library.  Allocates nothing -- the caller
    passes the outputs way[Q, max_iter+2, 2d] (float32) and nsteps[Q] (int32) -- and takes at most
    PLAN_STATIC_MAX_QUERIES queries per call.  Calls share the scratch: keep them on one stream."""
    Q = start_goal.shape[0]
    if way.numel() != Q * (max_iter + 2) * 2 * d or nsteps.numel() != Q:
        raise ValueError("plan_static: way must hold [Q, max_iter+2, 2d] floats and nsteps [Q] ints")
    _check(lib().pntf_plan_static(_ptr(params, name="params"), _ptr(B, name="B"), _ptr(start_goal, name="start_goal"),
                                  Q, d, float(eta), float(thr), int(max_iter), grad_flavour, _ptr(way),
                                  _ptr(nsteps, torch.int32), _stream()))
    return way, nsteps


def mesh_distance(triangles, points, offset=None, margin=None):
    """Unsigned distance of points [N,3] to the triangle soup triangles[M,3,3] (dataprocessing/speed_sampling_gpu.py:
    325-336); with offset / margin also speed = clip(distance, offset, margin) / margin (:415-419)."""
    dev = points.device
    M, N = triangles.shape[0], points.shape[0]
    dist = torch.empty(N, dtype=torch.float32, device=dev)
    speed = torch.empty(N, dtype=torch.float32, device=dev) if margin is not None else None
    ws = workspace(lib().pntf_mesh_workspace_bytes(M), dev)
    _check(lib().pntf_mesh_distance(_ptr(triangles, name="triangles"), M, _ptr(points, name="points"), N, _ptr(dist),
                                    float(offset or 0.0), float(margin or 0.0),
                                    _ptr(speed) if speed is not None else None, ws.data_ptr(), ws.numel(), _stream()))
    return (dist, speed) if speed is not None else dist


def adjoint_stats(E, N, d, budget=0, layer=-1):
    """{inf_nan, subnormal, nonzero, max_abs} of the fp16 adjoint tiles left in the cached workspace by the last
    loss(..., grad_out=..., prec=PREC_TC) call of the same shape (adjoints carry the library's scale).  layer = -1: all
    adjoint tiles, 0..12: Ybar_layer, 100 + l: the forward operand tile X_l (a -DPNTF_STASH_REGEN build does not
    stash X_0 / X_6 and raises for them), 200: the stashed encoder outputs Y5."""
    import struct
    dev = torch.device("cuda", torch.cuda.current_device())
    ws = workspace(workspace_bytes(OP_LOSS_BWD, E, N, d, PREC_TC, budget), dev)
    out = torch.zeros(4, dtype=torch.int64, device=dev)
    _check(lib().pntf_debug_adjoint_stats(ws.data_ptr(), ws.numel(), E, N, d, layer, out.data_ptr(), _stream()))
    v = out.cpu().tolist()
    return {"inf_nan": v[0], "subnormal": v[1], "nonzero": v[2],
            "max_abs": struct.unpack("f", struct.pack("I", v[3] & 0xffffffff))[0]}


def set_deterministic(on=True):
    """Run-to-run bit-identical parameter gradients in tensor-core mode (64-bit fixed-point accumulation)."""
    _check(lib().pntf_set_deterministic(1 if on else 0))


def tc_selftest(W, X, C):
    """out[3,128,128] = {W X, C + W X, X X^T} on the tensor cores (fp16 operands, fp32 accumulate)."""
    out = torch.empty(3, 128, 128, dtype=torch.float32, device=W.device)
    _check(lib().pntf_tc_selftest(_ptr(W), _ptr(X), _ptr(C), _ptr(out), _stream()))
    return out
