#!/usr/bin/env python
"""bench.py -- headline measurement of the P-NTFields hot path on B200.

Metric (BASELINE.json): Eikonal loss fwd+bwd point-pairs/s.  One "step" = one training step of
`Model.train()`'s hot loop (models/model_res_sigmoid_multi.py:1070-1080): Eikonal loss forward,
its backward (parameter gradients), gradient all-reduce when N > 1, and the AdamW update, over one
batch of synthetic start/goal pairs.  Workload at any N: BASELINE.json configs[1] "Gibson single
env 3D, 1M point pairs/batch, synthetic SDF speed" per GPU (weak scaling, configs[2]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--prec tc|fp32] [--d 3|6] [--pairs P]
    python bench.py --impl reference ...     # the reference's CPU path (oracle port) on the host cores

Prints ONE JSON line on rank 0.  `value` is device-resident throughput (inputs in HBM); `e2e` is the
same step through the reference-shaped public API (Model.Loss / backward / optimizer.step) with HOST
(pinned) inputs copied in and the loss read back inside the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "p-ntfields_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "eikonal_loss_fwd_bwd_point_pairs_per_sec"
UNIT = "pairs/s"
# algorithmic FLOPs per pair: 14 dense layers, 2 FLOP/MAC, all value/derivative
# columns; backward = 2x forward (dX and dW)
FWD_FLOP = {3: 14290176, 6: 27269376}
# DRAM bytes per tile of 128 rows (dram__bytes_read.sum + dram__bytes_write.sum of one launch / tiles of that launch)
# from the `ncu --set full` capture profiles/r01g_ncu_full_tc_d3.md (one 12 478-tile chunk, tensor-core path; the tile
# and stash formats are the same for d = 3 and d = 6); roofline.traffic = this x the tiles one launch of bench.py
# processes.  The loss kernels hold 12 pairs per tile at d = 3 and 8 at d = 6.
NCU_DRAM_BYTES_PER_TILE = {"tc_forward": (0.089265e9 + 9.351010e9) / 12478,
                           "tc_reverse": (8.077919e9 + 8.160507e9) / 12478,
                           "tc_dw": (16.357693e9 + 0.006265e9) / 12478}
LOSS_PAIRS_PER_TILE = {3: 12, 6: 8}

def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--prec", default=os.environ.get("PNTF_BENCH_PREC", "tc"), choices=["tc", "fp32"])
    ap.add_argument("--d", type=int, default=3, choices=[3, 6])
    ap.add_argument("--pairs", type=int, default=1 << 20, help="pairs per GPU per step")
    ap.add_argument("--envs", type=int, default=1)
    ap.add_argument("--cpu-pairs", type=int, default=10000, help="pairs per CPU-baseline chunk")
    ap.add_argument("--cpu-chunks", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--plan", action="store_true", help="(default on; kept for compatibility)")
    ap.add_argument("--no-plan", action="store_true", help="skip the path-integration metric in `extra`")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_burst": p["bf16_tflops"], "bf16_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm": p["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


# ---------------------------------------------------------------------------------------------------
# CPU leg: the oracle port of the reference's Loss + backward + AdamW on the host cores
# ---------------------------------------------------------------------------------------------------
def cpu_step_time(d, n_pairs, chunks, warm=1):
    """Seconds per chunk of n_pairs (median), threads used.  Chunk = the reference's own micro-batch
    (10 000 pairs, models/model_res_sigmoid_multi.py:996)."""
    import numpy as np
    import torch
    from oracle import pntf_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rs = np.random.RandomState(11)
    P = O.init_params(0)
    Bw = torch.from_numpy((0.5 * rs.normal(size=(1, d, 128))).astype(np.float32))
    c = torch.from_numpy(rs.uniform(-0.5, 0.5, size=(1, n_pairs, 2 * d)).astype(np.float32))
    s = torch.from_numpy(rs.uniform(0.1, 1.0, size=(1, n_pairs, 2)).astype(np.float32))
    flat = torch.cat([P[k].reshape(-1) for k in P])
    m = torch.zeros_like(flat)
    v = torch.zeros_like(flat)
    times = []
    for it in range(warm + chunks):
        t0 = time.perf_counter()
        loss, loss_n, diff, grads = O.loss_and_grads(P, c, s, Bw, 1.0, 1e-3, "multi")
        g = torch.cat([(grads[k] if grads[k] is not None else torch.zeros_like(P[k])).reshape(-1) for k in P])
        flat, m, v = O.adamw_step(flat, g, m, v, it + 1)
        float(loss_n)
        dt = time.perf_counter() - t0
        if it >= warm:
            times.append(dt)
    return statistics.median(times), threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_pairs
    sec, threads = cpu_step_time(args.d, n, max(args.steps, 1), warm=min(max(args.warmup, 1), 2))
    val = n / sec
    sample = "%d-pair chunk (the reference's micro-batch) of the same synthetic d=%d workload per step" % (n, args.d)
    out = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, "fp32"),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def workload_config(args, prec):
    return {"workload": "BASELINE configs[1]: Gibson single env d=%d training step (Eikonal loss fwd + bwd + AdamW), "
                        "%d point pairs per GPU per step, synthetic box/sphere SDF speed, random-init weights" % (args.d, args.pairs),
            "d": args.d, "pairs_per_gpu": args.pairs, "envs": args.envs, "loss_variant": "multi",
            "precision_mode": prec, "parallelism": "dp%d" % args.gpus,
            "l2": "256 MiB L2 flush write between timed steps (outside the per-step events)"}


# ---------------------------------------------------------------------------------------------------
# clocks sampling
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = "/tmp/pntf_clocks_%d_%d.csv" % (os.getpid(), index)

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons, pw = [], None, set(), []
        with open(self.path) as f:
            for line in f:
                parts = [x.strip() for x in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    mx = float(parts[2])
                    pw.append(float(parts[3]))
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        try:
            os.unlink(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": mx, "reasons": ["no samples"]}
        # "under load" = samples with power in the upper half of what was seen
        thr = 0.5 * (max(pw) + min(pw))
        load = [s for s, p in zip(sm, pw) if p >= thr] or sm
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(pw)}


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from pntf import capi
    from pntf import synth
    from pntf.model import Model_multi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback for the B200 arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    capi.lib()
    prec = args.prec
    d, E = args.d, args.envs
    n_total = args.pairs
    N = n_total // E
    n_total = N * E

    # ---- model + synthetic workload (seeded; rank-specific pairs, identical weights) ----
    torch.manual_seed(0)
    model = Model_multi("", "", d, E, device=str(dev))
    model.init_network()
    model.network.precision = prec
    net, opt = model.network, model.optimizer
    fp = net.flat()
    if world > 1:
        dist.broadcast(fp.flat, 0)
    B = synth.sample_B(d, dev, seed=1, E=E).contiguous()
    coords, speed = synth.sample_pairs(n_total, d, dev, seed=2 + rank, margin=0.5 / 12)
    alpha, beta, gamma = 1.0, 1.0, 1e-3
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    pv = capi.PREC_TC if prec == "tc" else capi.PREC_FP32
    from pntf.dist import GradAllReduce
    reducer = GradAllReduce(capi.PARAM_COUNT, dev)
    loss_out = torch.zeros(2, dtype=torch.float32, device=dev)
    diff = torch.empty(n_total, dtype=torch.float32, device=dev)

    def step_resident():
        fp.grad.zero_()
        capi.loss(capi.LOSS_MULTI, fp.flat, B, coords, speed, E, N, d, alpha, beta, gamma, grad_out=fp.grad, prec=pv,
                  diff=diff, loss_out=loss_out)
        reducer(fp.grad, loss_out)
        opt.step(grad_scale=1.0 / world)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_resident()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = capi.launch_count()
    capi.profile_begin()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in ev:
        flush.fill_(1)
        a.record()
        step_resident()
        b.record()
    barrier()
    prof = capi.profile_end()
    launches = capi.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * n_total * args.steps / (ms_total * 1e-3)
    final_loss = float(loss_out[1].item())

    # ---- e2e: reference-shaped public API with host (pinned) inputs ----
    e2e = None
    if not args.no_e2e:
        pts_h = coords.view(E, N, 2 * d).cpu().pin_memory()
        spd_h = speed.view(E, N, 2).cpu().pin_memory()
        B_h = B.cpu().pin_memory()

        def step_e2e():
            loss, loss_n, _ = model.Loss(pts_h, spd_h, B_h, beta, gamma)     # H2D inside
            loss.backward()
            if world > 1:
                reducer(fp.grad, torch.stack([loss.detach(), loss_n.detach()]))
            opt.step(grad_scale=1.0 / world)
            opt.zero_grad()
            return float(loss_n.detach())                                   # D2H read of the step's loss

        for _ in range(max(1, min(args.warmup, 2))):
            step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        barrier()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        e2e = {"value": world * n_total * args.steps / dt, "unit": UNIT,
               "h2d_bytes_per_step": int(pts_h.numel() * 4 + spd_h.numel() * 4 + B_h.numel() * 4), "d2h_bytes_per_step": 4,
               "ms_per_step": dt * 1e3 / args.steps, "api": "Model.Loss(host tensors) + loss.backward() + optimizer.step() + float(loss_n)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (CUDA events on the launching stream, inside the timed steps) ----
    pk = peaks()
    fl = FWD_FLOP[d]
    # algorithmic FLOPs each kernel class is responsible for, per pair
    class_flop = {"forward": fl, "reverse": fl, "dw": fl, "tc_forward": fl, "tc_reverse": fl, "tc_dw": fl}
    top = max(prof.items(), key=lambda kv: kv[1][0])
    top_name, (top_ms, top_cnt) = top
    step_ms_kernels = sum(v[0] for v in prof.values())
    roof = None
    if top_name in class_flop and top_ms > 0:
        flops = class_flop[top_name] * n_total * args.steps
        ach = flops / (top_ms * 1e-3) / 1e12
        peak = pk["bf16_sustained"]
        per_tile = NCU_DRAM_BYTES_PER_TILE.get(top_name) if prec == "tc" else None
        per_pair = per_tile / LOSS_PAIRS_PER_TILE[d] if per_tile else None
        pairs_per_launch = n_total * args.steps / max(top_cnt, 1)
        traffic = per_pair * pairs_per_launch if per_pair else None
        roof = {"bound": "tensor", "kernel": top_name, "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                "traffic": traffic,
                "traffic_source": "profiles/r01g_ncu_full_tc_d3.md (ncu --set full, bytes per 128-row tile x tiles per launch)" if per_pair else None,
                "hbm_GBps": (per_pair * n_total * args.steps / (top_ms * 1e-3) / 1e9) if per_pair else None,
                "hbm_peak_GBps": pk["hbm"],
                "hbm_frac": (per_pair * n_total * args.steps / (top_ms * 1e-3) / 1e9 / pk["hbm"]) if per_pair else None,
                "note": "the training kernels move 3.37 MB of activation stash per 128-row tile (12 pairs at d=3, 8 at d=6) through "
                        "HBM; see DESIGN.md section 7 (forward 86 % of the 3.89 TB/s write-only peak, dW at the read peak)", "peak_source": pk["source"] + ", sustained bf16 (kernel timed inside a long step)",
                "launch_ms_avg": top_ms / max(top_cnt, 1), "launches": top_cnt,
                "share_of_step": top_ms / max(step_ms_kernels, 1e-9),
                "alg_flop_per_pair": class_flop[top_name],
                "whole_step_tflops": (3 * fl) * n_total * args.steps / (ms_total * 1e-3) / 1e12}
    kernels = {k: {"ms_per_step": v[0] / args.steps, "launches_per_step": v[1] / args.steps} for k, v in prof.items() if v[1]}

    cpu = None
    if not args.no_cpu_baseline:
        sec, threads = cpu_step_time(d, args.cpu_pairs, args.cpu_chunks)
        cpu = {"value": args.cpu_pairs / sec, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "%d chunks of %d pairs (the reference's micro-batch) of the same workload, oracle port of Model.Loss + backward + AdamW, torch CPU fp32"
                         % (args.cpu_chunks, args.cpu_pairs)}

    extra = {}
    if not args.no_plan:
        extra["plan"] = plan_metric(model, B, d, dev, cpu=not args.no_cpu_baseline)
        extra["fields"] = fields_metric(fp.flat, B, coords, d, pv)
        if d == 3:
            extra["speed_sampling"] = speed_sampling_metric(dev, cpu=not args.no_cpu_baseline)

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f16 tensor-core operands, f32 accumulate/master" if prec == "tc" else "f32",
           "data": "synthetic", "config": workload_config(args, prec), "clocks": clocks, "e2e": e2e,
           "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "kernels": kernels,
           "final_loss_n": final_loss}
    if extra:
        out["extra"] = extra
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def speed_sampling_metric(dev, cpu=True, N=1 << 20, levels=6):
    """Secondary metric (dataprocessing/speed_sampling_gpu.py:325-421): unsigned point-to-mesh
    distance queries/s for 1 Mi points against a synthetic scene of three boxes subdivided to 147 456 triangles, and
    sampled pairs/s of the whole rejection sampler (random draws, two distance queries, filtering) in the Gibson band."""
    import torch
    from pntf import capi, synth, speed_sampling as SS
    t = synth.box_scene_mesh(dev, levels)                                          # shuffled: mesh files are not sorted
    g = torch.Generator(device=dev).manual_seed(11)
    pts = torch.rand(N, 3, device=dev, generator=g) - 0.5
    mesh = SS.ObstacleMesh(t.unsqueeze(0))

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps * 1e-3

    order = SS._morton_order(pts)
    ps = pts[order].contiguous()
    sec_kernel = timed(lambda: capi.mesh_distance(mesh.triangles, ps))
    sec_api = timed(lambda: mesh.distance(pts))
    sec_unsorted = timed(lambda: capi.mesh_distance(t, pts), reps=1)
    offset, margin = SS.band_for_task("gibson")
    want = 100000
    torch.manual_seed(5)
    X, Y = [], []
    sec_sampler = timed(lambda: SS.point_append_list(X, Y, mesh, want, 3, offset, margin), reps=1)
    res = {"triangles": int(t.shape[0]), "points": N,
           "distance_queries_per_s": N / sec_kernel, "distance_queries_per_s_with_host_sort": N / sec_api,
           "distance_queries_per_s_unsorted_inputs": N / sec_unsorted,
           "brute_force_equivalent_pair_tests_per_s": N * float(t.shape[0]) / sec_kernel,
           "sampler_pairs_per_s": want / sec_sampler, "sampler_pairs": want, "band": [offset, margin]}
    if cpu:
        from oracle import mesh_oracle as MO                                       # CPU leg only
        n = 256
        p = pts[:n].cpu().numpy()
        tri = t.cpu().numpy()
        t0 = time.perf_counter()
        MO.point_obstacle_distance(p, tri)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"distance_queries_per_s": n / dt, "kind": "port", "cores": 1,
                               "sample": "%d of the points, brute force over all triangles in numpy float64" % n}
    return res


def fields_metric(flat, B, coords, d, prec):
    """Secondary metric: inference fields of SURVEY 8a row I1 (Model.TravelTimes / Speed / Speed2 / Gradient) and the three
    forward modes on the step's pairs, pairs/s with inputs resident (CUDA events, best of 3)."""
    import torch
    from pntf import capi
    n = coords.shape[0]
    Bq = B[:1].contiguous()
    res = {"pairs": n}

    def best(f):
        f()
        torch.cuda.synchronize()
        t = 1e9
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            f()
            b.record()
            torch.cuda.synchronize()
            t = min(t, a.elapsed_time(b))
        return n / (t * 1e-3)

    for name, kind in (("travel_times", 0), ("speed", 1), ("speed2", 2), ("direction", 3)):
        res[name + "_pairs_per_s"] = best(lambda: capi.field(kind, flat, Bq, coords, 1, n, d, gamma=1e-3, prec=prec))
    for name, mode in (("out", capi.MODE_VALUE), ("out_grad", capi.MODE_GRAD), ("out_laplace", capi.MODE_LAPLACE)):
        res[name + "_pairs_per_s"] = best(lambda: capi.forward(mode, flat, Bq, coords, 1, n, d, prec=prec))
    return res


def plan_metric(model, B, d, dev, Q=65536, cpu=True):
    """Secondary metric: batched bidirectional path integration, paths/s (BASELINE configs[4]): 65 536 start/goal
    queries ~ U[-0.4,0.4]^d, reference step / threshold / cap (test/gib_plan.py:74-86, test/arm_plan.py:140-152), on
    freshly initialised weights (the trained weights of the timed steps are left alone)."""
    import torch
    from pntf import capi, synth
    from pntf.model import Model_multi
    q = synth.sample_queries(Q, d, dev)
    eta, thr, cap = (0.03, 0.06, 500) if d == 3 else (0.015, 0.03, 300)
    torch.manual_seed(0)
    fresh = Model_multi("", "", d, 1, device=str(dev))
    fresh.init_network()
    flat = fresh.network.flat().flat
    Bq = B[:1].contiguous()
    res = {"queries": Q, "eta": eta, "thr": thr, "max_iter": cap}
    for mode in (("compat", "exact") if d == 3 else ("exact",)):
        fl = capi.GRAD_COMPAT if mode == "compat" else capi.GRAD_EXACT
        for prec_name, prec in (("tc", capi.PREC_TC), ("fp32_simt", capi.PREC_FP32)):
            if prec_name == "fp32_simt" and mode != "exact":
                continue
            if prec_name == "tc":      # warm-up at full size (the waypoint buffer comes from torch's caching allocator)
                capi.plan(flat, Bq, q, d, eta, thr, cap, fl, prec=prec)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            way, nst = capi.plan(flat, Bq, q, d, eta, thr, cap, fl, prec=prec)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            steps = float(nst.float().sum().item())
            res["%s_%s" % (mode, prec_name)] = {
                "paths_per_s": Q / (ms * 1e-3), "path_steps_per_s": steps / (ms * 1e-3), "mean_steps": steps / Q,
                "ms": ms, "finite": bool(torch.isfinite(way).all().item()),
                "alg_waypoint_GBps": steps * 2 * d * 4 / (ms * 1e-3) / 1e9,
                "written_GBps": way.numel() * 4 / (ms * 1e-3) / 1e9,
                "alg_tflops": steps * 2621952 / (ms * 1e-3) / 1e12 if d == 3 else None}
    if cpu:
        res["cpu_baseline"] = plan_cpu_baseline(q[:256].cpu(), Bq[0].cpu(), d, eta, thr, cap)
    return res


def plan_cpu_baseline(q, Bw, d, eta, thr, cap):
    """The reference plan loop (batched restatement of test/gib_plan.py:53-86 with per-row termination, oracle port)
    on the host cores: 256 of the same queries, exact gradient, fp32."""
    import torch
    from oracle import pntf_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    P = O.init_params(0)
    with torch.no_grad():
        O.plan(P, q[:16].clone(), Bw, eta, thr, cap, "exact")          # warm-up
        t0 = time.perf_counter()
        way, nst = O.plan(P, q.clone(), Bw, eta, thr, cap, "exact")
        dt = time.perf_counter() - t0
    steps = float(nst.sum())
    return {"paths_per_s": q.shape[0] / dt, "path_steps_per_s": steps / dt, "mean_steps": steps / q.shape[0],
            "cores": threads, "kind": "port", "sample": "256 of the 65 536 queries, batched with per-row termination"}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
