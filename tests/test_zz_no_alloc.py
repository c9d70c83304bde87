"""pntf_plan_static: the allocation-free variant of the tensor-core path integration (scratch in a file-scope __device__
array of the library, include/pntf_b200.h).  No counterpart in the reference; checked against the existing entry point
pntf_plan(prec = PNTF_PREC_TC) through the C-ABI, on two calls in a row that share the scratch.  Collected last on
purpose."""
import ctypes

import numpy as np
import pytest
import torch

DEV = "cuda:0"


def _lib():
    from pntf import capi
    return capi, capi.lib()


def test_static_plan_rejects_more_queries_than_its_scratch_holds():
    # argument checking happens before any CUDA call: runs without a GPU
    capi, L = _lib()
    dummy = ctypes.c_void_p(256)
    Q = capi.PLAN_STATIC_MAX_QUERIES + 1
    rc = L.pntf_plan_static(dummy, dummy, dummy, Q, 3, 0.03, 0.06, 10, 0, dummy, dummy, None)
    assert rc != 0
    msg = L.pntf_last_error().decode()
    assert "PNTF_PLAN_STATIC_MAX_QUERIES" in msg and str(Q) in msg
    assert L.pntf_plan_static(dummy, dummy, dummy, 8, 9, 0.03, 0.06, 10, 0, dummy, dummy, None) != 0      # d = 9
    assert L.pntf_plan_static(dummy, dummy, dummy, 8, 3, 0.03, 0.06, 10, 0, None, dummy, None) != 0       # no output


def _case(seed, d, Q):
    from oracle import pntf_oracle as O
    rs = np.random.RandomState(seed)
    P = O.init_params(8 + seed)
    flat = torch.cat([P[k].reshape(-1) for k in P]).to(DEV).contiguous()
    Bw = torch.from_numpy((0.5 * rs.normal(size=(1, d, 128))).astype(np.float32)).to(DEV)
    sg = rs.uniform(-0.4, 0.4, size=(Q, 2 * d)).astype(np.float32)
    sg[3, d:] = sg[3, :d] + 0.005          # already within thr: zero steps
    return flat, Bw, torch.from_numpy(sg).to(DEV)


@pytest.mark.gpu
def test_static_plan_matches_the_workspace_plan_on_two_calls_in_a_row():
    capi, L = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    # the second call is smaller, of the other dimension and of other weights: anything it picked up from the
    # first call's scratch (weight image, counters, hand-out order) would show
    for seed, d, Q, cap, eta, thr in [(1, 3, 3001, 40, 0.03, 0.06), (2, 6, 257, 25, 0.015, 0.03), (3, 3, 19, 12, 0.03, 0.06)]:
        flat, Bw, sg = _case(seed, d, Q)
        ref_way, ref_n = capi.plan(flat, Bw, sg, d, eta, thr, cap, capi.GRAD_EXACT, prec=capi.PREC_TC)
        way = torch.full((Q, cap + 2, 2 * d), float("nan"), dtype=torch.float32, device=DEV)
        nst = torch.full((Q,), -1, dtype=torch.int32, device=DEV)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = L.pntf_plan_static(flat.data_ptr(), Bw.data_ptr(), sg.data_ptr(), Q, d, eta, thr, cap, capi.GRAD_EXACT,
                                way.data_ptr(), nst.data_ptr(), stream)
        assert rc == 0, L.pntf_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        n, rn = nst.cpu().numpy(), ref_n.cpu().numpy()
        w, rw = way.cpu().numpy(), ref_way.cpu().numpy()
        assert np.isfinite(w).all() and n.min() >= 0 and n[3] == 0
        # tolerances of tests/test_gpu_tc.py::test_tc_plan_vs_oracle: a query at rounding distance of thr may take one
        # step more or less; waypoints within 1e-3 of the workspace scale (unit box)
        same = n == rn
        assert same.mean() > 0.97 and np.abs(n - rn).max() <= 1
        assert np.abs(w[same] - rw[same]).max() < 1e-3
        for q in range(0, Q, max(1, Q // 50)):
            assert np.all(w[q, n[q]:] == w[q, n[q]])


@pytest.mark.gpu
def test_static_plan_at_its_limit_and_through_the_wrapper():
    capi, L = _lib()
    d, cap = 3, 20
    Q = capi.PLAN_STATIC_MAX_QUERIES
    flat, Bw, sg = _case(4, d, Q)
    ref_way, ref_n = capi.plan(flat, Bw, sg, d, 0.03, 0.06, cap, capi.GRAD_EXACT, prec=capi.PREC_TC)
    way = torch.empty(Q, cap + 2, 2 * d, dtype=torch.float32, device=DEV)
    nst = torch.empty(Q, dtype=torch.int32, device=DEV)
    for _ in range(2):
        way.fill_(float("nan"))
        capi.plan_static(flat, Bw, sg, d, 0.03, 0.06, cap, way, nst)
        n, rn = nst.cpu().numpy(), ref_n.cpu().numpy()
        same = n == rn
        assert same.mean() > 0.97 and np.abs(n - rn).max() <= 1
        assert (way[torch.from_numpy(same).to(DEV)] - ref_way[torch.from_numpy(same).to(DEV)]).abs().max().item() < 1e-3
    with pytest.raises(RuntimeError, match="PNTF_PLAN_STATIC_MAX_QUERIES"):
        sg2 = torch.zeros(Q + 1, 2 * d, device=DEV)
        capi.plan_static(flat, Bw, sg2, d, 0.03, 0.06, 1, torch.empty(Q + 1, 3, 2 * d, device=DEV),
                         torch.empty(Q + 1, dtype=torch.int32, device=DEV))
