#!/usr/bin/env python
"""bench.py -- headline measurement of the P-NTFields hot path on B200.

Metric (BASELINE.json): Eikonal loss fwd+bwd point-pairs/s.  One "step" = one training step of
`Model.train()`'s hot loop (models/model_res_sigmoid_multi.py:1070-1080): Eikonal loss forward,
its backward (parameter gradients), gradient all-reduce when N > 1, and the AdamW update, over one
batch of synthetic start/goal pairs.  Workload at any N: BASELINE.json configs[1] "Gibson single
env 3D, 1M point pairs/batch, synthetic SDF speed" per GPU (weak scaling, configs[2]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--prec tc|fp32] [--d 3|6] [--pairs P]
    python bench.py --impl reference ...     # the reference's own CPU code (oracle/_ref) on the host cores

Prints ONE JSON line on rank 0.  `value` is device-resident throughput (inputs in HBM); `e2e` is the
same step through the reference-shaped public API (Model.Loss / backward / optimizer.step) with HOST
(pinned) inputs copied in and the loss read back inside the timed region.  `extra` carries the other
BASELINE configs measured in the same run: `test_box` (configs[0]), `strong` (configs[2], strong scaling,
E = 2), `d6` (configs[3] + the UR5 half of configs[4]), `plan` (configs[4], sharded over the ranks),
`small_batch` (the reference's own 2 x 10 000-pair step through Model.train's loop), `fields`,
`speed_sampling`.
"""
import argparse
import glob
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
PLAN_FLOP_PER_STEP = 2621952              # value forward + reverse sweep per query-step, d = 3
LOSS_PAIRS_PER_TILE = {3: 12, 6: 8}
MARGIN = {"gibson": 0.5 / 12, "arm": 0.5 / 12, "test": 0.5 / 5}     # dataprocessing/speed_sampling_gpu.py:467-475


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
    ap.add_argument("--no-plan", action="store_true", help="skip every `extra` metric")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_burst": p["bf16_tflops"], "bf16_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm": p["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def dram_bytes_per_tile():
    """DRAM bytes per 128-row tile and kernel class from the newest committed ncu capture
    (profiles/*dram_bytes_per_tile.json, written by scripts/ncu_traffic.py from `ncu --set full`)."""
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*dram_bytes_per_tile.json")))
    if not files:
        return {}, None
    with open(files[-1]) as f:
        j = json.load(f)
    return j.get("bytes_per_tile", {}), os.path.relpath(files[-1], ROOT)


# ---------------------------------------------------------------------------------------------------
# CPU leg: the reference's own Model.Loss + backward + AdamW (oracle/_ref) on the host cores; the oracle
# port only when oracle/_ref did not travel
# ---------------------------------------------------------------------------------------------------
def cpu_step_time(d, n_pairs, chunks, warm=1, margin=MARGIN["gibson"]):
    """Seconds per chunk of n_pairs (median), threads used, kind.  Chunk = the reference's own micro-batch
    (10 000 pairs, models/model_res_sigmoid_multi.py:996).  d = 3: multi-environment model (E = 1), d = 6: the
    single-environment model (UR5), as the reference trains them."""
    import torch
    from oracle import make_ref
    from pntf import synth
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    c, s = synth.sample_pairs(n_pairs, d, torch.device("cpu"), seed=2, margin=margin)
    times = []
    if make_ref.available():
        kind = "reference"
        S, M = make_ref.load()
        torch.manual_seed(0)
        if d == 3:
            model = M.Model("/tmp", "/tmp", d, 1, device="cpu")
            model.network = M.NN("cpu", d)
            B = synth.sample_B(d, "cpu", seed=1, E=1)
            call = lambda: model.Loss(c.view(1, n_pairs, 2 * d), s.view(1, n_pairs, 2), B, 1.0, 1e-3)
        else:
            model = S.Model("/tmp", "/tmp", d, device="cpu")
            model.network = S.NN("cpu", d, 0.5 * torch.randn(128, d))
            call = lambda: model.Loss(c, s, 1.0, 1e-3)
        model.network.apply(model.network.init_weights)
        opt = torch.optim.AdamW(model.network.parameters(), lr=1e-3, weight_decay=0.1)
        for it in range(warm + chunks):
            t0 = time.perf_counter()
            loss, loss_n, _ = call()
            loss.backward()
            opt.step()
            opt.zero_grad()
            float(loss_n)
            dt = time.perf_counter() - t0
            if it >= warm:
                times.append(dt)
    else:
        kind = "port"
        from oracle import pntf_oracle as O
        P = O.init_params(0)
        Bw = synth.sample_B(d, "cpu", seed=1, E=1)
        flat = torch.cat([P[k].reshape(-1) for k in P])
        m, v = torch.zeros_like(flat), torch.zeros_like(flat)
        variant = "multi" if d == 3 else "single"
        for it in range(warm + chunks):
            t0 = time.perf_counter()
            if variant == "multi":
                loss, loss_n, diff, grads = O.loss_and_grads(P, c.unsqueeze(0), s.unsqueeze(0), Bw, 1.0, 1e-3, "multi")
            else:
                loss, loss_n, diff, grads = O.loss_and_grads(P, c, s, Bw[0], 1.0, 1e-3, "single")
            g = torch.cat([(grads[k] if grads[k] is not None else torch.zeros_like(P[k])).reshape(-1) for k in P])
            flat, m, v = O.adamw_step(flat, g, m, v, it + 1)
            float(loss_n)
            dt = time.perf_counter() - t0
            if it >= warm:
                times.append(dt)
    return statistics.median(times), threads, kind


def cpu_baseline(d, n_pairs, chunks, margin=MARGIN["gibson"]):
    sec, threads, kind = cpu_step_time(d, n_pairs, chunks, margin=margin)
    what = ("the reference's own Model.Loss + loss.backward() + torch.optim.AdamW.step (oracle/_ref, unmodified "
            "models/model_res_sigmoid%s.py)" % ("_multi" if d == 3 else "")) if kind == "reference" else \
        "oracle port of Model.Loss + backward + AdamW (oracle/_ref absent)"
    return {"value": n_pairs / sec, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": "%d chunks of %d pairs (the reference's micro-batch) of the same synthetic d=%d workload, %s, torch CPU fp32"
                      % (chunks, n_pairs, d, what)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_pairs
    sec, threads, kind = cpu_step_time(args.d, n, max(args.steps, 1), warm=min(max(args.warmup, 1), 2))
    val = n / sec
    sample = "%d-pair chunk (the reference's micro-batch) of the same synthetic d=%d workload per step" % (n, args.d)
    out = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, "fp32"),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def workload_config(args, prec):
    return {"workload": "BASELINE configs[1]: Gibson single env d=%d training step (Eikonal loss fwd + bwd + AdamW), "
                        "%d point pairs per GPU per step, synthetic box/sphere SDF speed, random-init weights" % (args.d, args.pairs),
            "d": args.d, "pairs_per_gpu": args.pairs, "envs": args.envs, "loss_variant": "multi" if args.d == 3 else "single",
            "precision_mode": prec, "parallelism": "dp%d" % args.gpus,
            "l2": "256 MiB L2 flush write before every timed step of `value` and of `e2e` (outside the per-step timing)"}


# ---------------------------------------------------------------------------------------------------
# clocks sampling
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    # power.draw is a ~1 s average and under-reads a sub-second timed region (it shows 550-620 W where the board sits at its
    # 1000 W limit: scripts/power_probe.sh, profiles/r02x_power_probe.txt); it is still what selects the samples "under
    # load" (it ramps smoothly).  power.draw.instant, the last column, is reported beside it (spiky: 1000-1200 W peaks).
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,"
         "power.draw.instant")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = "/tmp/pntf_clocks_%d_%d.csv" % (os.getpid(), index)

    def start(self):
        try:
            # a driver without the instantaneous field would reject the whole query: probe it once, outside the timed region
            probe = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=power.draw.instant", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=20)
            if probe.returncode != 0 or "not a valid field" in (probe.stdout + probe.stderr).lower():
                self.Q = self.Q.rsplit(",", 1)[0]
        except (OSError, subprocess.SubprocessError):
            pass
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=self.f, stderr=subprocess.DEVNULL)
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
        sm, mx, reasons, pw, pwi = [], None, set(), [], []
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
                try:
                    pwi.append(float(parts[9]))
                except (ValueError, IndexError):
                    pass
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
        res = {"sm_mhz": statistics.median(load), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
               "power_w_max": max(pw)}
        if pwi:
            res["power_w_instant_max"] = max(pwi)
        return res


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
class Ctx:
    pass


def make_training(ctx, d, E, n_total, margin, seed=2):
    """Model + resident synthetic batch + the step closure of one configuration."""
    import torch
    from pntf import capi, synth
    from pntf.dist import GradAllReduce
    from pntf.model import Model_multi, Model_single
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    N = n_total // E
    n_total = N * E
    torch.manual_seed(0)
    if d == 3:
        model = Model_multi("", "", d, E, device=str(dev))
        model.init_network()
        B = synth.sample_B(d, dev, seed=1, E=E).contiguous()
        variant = capi.LOSS_MULTI
    else:       # UR5: the single-environment model (square roots in the residual, no |B|^2 term)
        model = Model_single("", "", d, device=str(dev))
        model.init_network(B=synth.sample_B(d, "cpu", seed=1, E=1)[0].T.contiguous())
        B = model.network.B.to(torch.float32).contiguous().view(1, d, 128)
        variant = capi.LOSS_SINGLE
    model.network.precision = ctx.prec
    fp, opt = model.network.flat(), model.optimizer
    if world > 1:
        import torch.distributed as dist
        dist.broadcast(fp.flat, 0)
    coords, speed = synth.sample_pairs(n_total, d, dev, seed=seed + rank, margin=margin)
    reducer = GradAllReduce(capi.PARAM_COUNT, dev)
    diff = torch.empty(n_total, dtype=torch.float32, device=dev)
    pv = capi.PREC_TC if ctx.prec == "tc" else capi.PREC_FP32

    def step():
        fp.grad.zero_()
        capi.loss(variant, fp.flat, B, coords, speed, E, N, d, 1.0, 1.0, 1e-3, grad_out=fp.grad, prec=pv, diff=diff,
                  loss_out=fp.loss_slot)
        reducer.reduce_flat(fp)
        opt.step(grad_scale=1.0 / world)

    t = Ctx()
    t.model, t.fp, t.opt, t.B, t.coords, t.speed, t.step, t.E, t.N, t.n_total, t.d, t.variant, t.reducer, t.pv = \
        model, fp, opt, B, coords, speed, step, E, N, n_total, d, variant, reducer, pv
    return t


def barrier(ctx):
    import torch
    torch.cuda.synchronize()
    if ctx.world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def time_steps(ctx, step, steps, warmup, flush=True):
    """Total ms of `steps` calls (CUDA events around each, L2 flush before each, max over ranks)."""
    import torch
    for _ in range(warmup):
        step()
    barrier(ctx)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        if flush:
            ctx.flush.fill_(1)
        a.record()
        step()
        b.record()
    barrier(ctx)
    ms = sum(a.elapsed_time(b) for a, b in ev)
    return max_over_ranks(ctx, ms)


def max_over_ranks(ctx, x):
    import torch
    if ctx.world == 1:
        return float(x)
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=ctx.dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def kernel_profile(ctx, step, steps):
    """Per-kernel-class device time of `steps` steps (CUDA events inside the library, separate pass)."""
    from pntf import capi
    barrier(ctx)
    capi.profile_begin()
    for _ in range(steps):
        ctx.flush.fill_(1)
        step()
    barrier(ctx)
    return capi.profile_end()


def roofline_of(ctx, prof, d, n_total, steps, ms_total):
    pk = peaks()
    fl = FWD_FLOP[d]
    # algorithmic FLOPs each kernel class is responsible for, per pair (the fused backward launch does dX and dW)
    class_flop = {"forward": fl, "reverse": fl, "dw": fl, "tc_forward": fl, "tc_reverse": fl, "tc_dw": fl, "tc_backward": 2 * fl}
    top_name, (top_ms, top_cnt) = max(prof.items(), key=lambda kv: kv[1][0])
    if top_name not in class_flop or top_ms <= 0:
        return None
    step_ms_kernels = sum(v[0] for v in prof.values())
    ach = class_flop[top_name] * n_total * steps / (top_ms * 1e-3) / 1e12
    peak = pk["bf16_sustained"]
    table, src = dram_bytes_per_tile()
    per_tile = table.get(top_name) if ctx.prec == "tc" else None
    per_pair = per_tile / LOSS_PAIRS_PER_TILE[d] if per_tile else None
    pairs_per_launch = n_total * steps / max(top_cnt, 1)
    alg_bytes_pair = 2 * d * 4 + 8 + 4
    return {"bound": "tensor", "limiter": "hbm (the activation stash) at full clocks; board power (energy per step) and the serial MMA->epilogue chain under a power cap (DESIGN.md section 7b)", "kernel": top_name,
            "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
            "traffic": per_pair * pairs_per_launch if per_pair else None, "traffic_source": src if per_pair else None,
            "dram_bytes_per_pair": per_pair, "algorithmic_bytes_per_pair": alg_bytes_pair,
            "hbm_GBps": (per_pair * n_total * steps / (top_ms * 1e-3) / 1e9) if per_pair else None,
            "hbm_peak_GBps": pk["hbm"],
            "hbm_frac": (per_pair * n_total * steps / (top_ms * 1e-3) / 1e9 / pk["hbm"]) if per_pair else None,
            "peak_source": pk["source"] + ", sustained bf16 (kernel timed inside a long step)",
            "launch_ms_avg": top_ms / max(top_cnt, 1), "launches": top_cnt,
            "share_of_step": top_ms / max(step_ms_kernels, 1e-9), "alg_flop_per_pair": class_flop[top_name],
            "whole_step_tflops": (3 * fl) * n_total * steps / (ms_total * 1e-3) / 1e12,
            "whole_step_frac_of_sustained": (3 * fl) * n_total * steps / (ms_total * 1e-3) / 1e12 / peak}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from pntf import capi

    ctx = Ctx()
    ctx.world = int(os.environ.get("WORLD_SIZE", "1"))
    ctx.rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback for the B200 arm")
    torch.cuda.set_device(local)
    ctx.dev = torch.device("cuda", local)
    if ctx.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=ctx.dev)
    capi.lib()
    ctx.prec = args.prec
    ctx.flush = torch.empty(256 << 20, dtype=torch.uint8, device=ctx.dev)
    world, rank, d = ctx.world, ctx.rank, args.d

    # ---- main workload (BASELINE configs[1]; weak scaling: the same batch size on every rank) ----
    t = make_training(ctx, d, args.envs, args.pairs, MARGIN["gibson"])
    for _ in range(args.warmup):
        t.step()
    barrier(ctx)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = capi.launch_count()
    ms_total = time_steps(ctx, t.step, args.steps, 0)
    launches = capi.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms_total / args.steps
    value = world * t.n_total * args.steps / (ms_total * 1e-3)
    prof = kernel_profile(ctx, t.step, args.steps)
    final_loss = float(t.fp.loss_slot[1].item())

    # ---- e2e: reference-shaped public API with host (pinned) inputs ----
    e2e = None
    if not args.no_e2e:
        e2e = e2e_metric(ctx, t, args)

    extra = {}
    if not args.no_plan:
        cpu = not args.no_cpu_baseline
        # configs[4] (d of the run), queries sharded over the ranks, no collective
        extra["plan"] = plan_metric(ctx, t.B, d, cpu=cpu and rank == 0)
        # configs[2], strong scaling: 1 Mi pairs GLOBAL per step, two environments with their own B
        if d == 3:
            extra["strong"] = strong_metric(ctx, args)
        if world == 1:
            extra["fields"] = fields_metric(t.fp.flat, t.B, t.coords, d, t.pv)
            if d == 3:
                extra["test_box"] = test_box_metric(ctx, args)
                extra["small_batch"] = small_batch_metric(ctx)
                extra["speed_sampling"] = speed_sampling_metric(ctx.dev, cpu=cpu)
                del t
                torch.cuda.empty_cache()
                extra["d6"] = d6_metric(ctx, args, cpu=cpu)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    roof = roofline_of(ctx, prof, d, args.pairs // args.envs * args.envs, args.steps, ms_total)
    kernels = {k: {"ms_per_step": v[0] / args.steps, "launches_per_step": v[1] / args.steps} for k, v in prof.items() if v[1]}
    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline(d, args.cpu_pairs, args.cpu_chunks)

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f16 tensor-core operands, f32 accumulate/master" if args.prec == "tc" else "f32",
           "data": "synthetic", "config": workload_config(args, args.prec), "clocks": clocks, "e2e": e2e,
           "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "kernels": kernels,
           "kernels_note": "per-class times from a separate pass of the same steps with CUDA events inside the library; "
                           "`value` is timed without them",
           "final_loss_n": final_loss}
    if extra:
        out["extra"] = extra
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def e2e_metric(ctx, t, args):
    import torch
    world, d, E, N = ctx.world, t.d, t.E, t.N
    model, fp, opt = t.model, t.fp, t.opt
    pts_h = (t.coords.view(E, N, 2 * d) if d == 3 else t.coords).cpu().pin_memory()
    spd_h = (t.speed.view(E, N, 2) if d == 3 else t.speed).cpu().pin_memory()
    B_h = t.B.cpu().pin_memory()

    def step_e2e():
        if d == 3:
            loss, loss_n, _ = model.Loss(pts_h, spd_h, B_h, 1.0, 1e-3)     # H2D inside
        else:
            loss, loss_n, _ = model.Loss(pts_h, spd_h, 1.0, 1e-3)
        loss.backward()
        if world > 1:
            fp.loss_slot.copy_(torch.stack([loss.detach(), loss_n.detach()]))
            t.reducer.reduce_flat(fp)
        opt.step(grad_scale=1.0 / world)
        opt.zero_grad()
        return float(loss_n.detach())                                   # D2H read of the step's loss

    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    barrier(ctx)
    dt = 0.0
    for _ in range(args.steps):
        ctx.flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step_e2e()
        torch.cuda.synchronize()
        dt += time.perf_counter() - t0
    barrier(ctx)
    dt = max_over_ranks(ctx, dt)
    return {"value": world * t.n_total * args.steps / dt, "unit": UNIT,
            "h2d_bytes_per_step": int(pts_h.numel() * 4 + spd_h.numel() * 4 + B_h.numel() * 4), "d2h_bytes_per_step": 4,
            "ms_per_step": dt * 1e3 / args.steps,
            "api": "Model.Loss(host tensors) + loss.backward() + optimizer.step() + optimizer.zero_grad() + float(loss_n); "
                   "wall clock around each step, synchronised on both sides"}


def strong_metric(ctx, args):
    """BASELINE configs[2], strong scaling: ONE global batch of 1 Mi pairs per step, E = 2
    environments with their own Fourier matrices, split by pair range inside each environment over the ranks."""
    import torch
    world = ctx.world
    glob_pairs = 1 << 20
    E = 2
    per_rank = glob_pairs // world
    t = make_training(ctx, 3, E, per_rank, MARGIN["gibson"], seed=40)
    steps = max(3, args.steps)
    ms = time_steps(ctx, t.step, steps, 3)
    res = {"workload": "BASELINE configs[2]: 1 048 576 pairs GLOBAL per step, E=2 environments, split evenly over %d rank(s)" % world,
           "scaling": "strong", "global_pairs": per_rank * world, "pairs_per_gpu": per_rank, "envs": E, "n_gpus": world,
           "ms_per_step": ms / steps, "value": per_rank * world * steps / (ms * 1e-3), "unit": UNIT}
    del t
    torch.cuda.empty_cache()
    return res


def test_box_metric(ctx, args):
    """BASELINE configs[0]: test box environment, d = 3, 100 000 pairs, margin 0.5/5 -- the training
    step with E = 1 x 100 000 and with E = 2 x 50 000; its CPU side is the line's cpu_baseline (same d = 3 model)."""
    import torch
    res = {"workload": "BASELINE configs[0]: test box env (configs/test.txt margin 0.5/5), d=3, Eikonal loss fwd+bwd+AdamW on 100 000 pairs"}
    for E in (1, 2):
        t = make_training(ctx, 3, E, 100000, MARGIN["test"], seed=50)
        ms = time_steps(ctx, t.step, 5, 3)
        res["E%d" % E] = {"pairs": t.n_total, "ms_per_step": ms / 5, "value": t.n_total * 5 / (ms * 1e-3), "unit": UNIT,
                          "loss_n": float(t.fp.loss_slot[1].item())}
        del t
    torch.cuda.empty_cache()
    return res


def d6_metric(ctx, args, cpu=True):
    """BASELINE configs[3] (UR5 6-DOF, d = 6, single-environment model, margin 0.5/12) training step with its own roofline
    and CPU baseline, and the UR5 half of configs[4] (path integration, eta 0.015, thr 0.03, cap 300)."""
    import torch
    t = make_training(ctx, 6, 1, args.pairs, MARGIN["arm"])
    steps = 3
    ms = time_steps(ctx, t.step, steps, 3)
    prof = kernel_profile(ctx, t.step, steps)
    res = {"workload": "BASELINE configs[3]: UR5 6-DOF C-space (configs/arm.txt), d=6, single-env model, %d pairs per step, "
                       "fwd + grad tau + Laplacian + loss + bwd + AdamW" % t.n_total,
           "value": t.n_total * steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
           "roofline": roofline_of(ctx, prof, 6, t.n_total, steps, ms),
           "kernels": {k: {"ms_per_step": v[0] / steps} for k, v in prof.items() if v[1]},
           "loss_n": float(t.fp.loss_slot[1].item())}
    B = t.B
    del t
    torch.cuda.empty_cache()
    if cpu:
        res["cpu_baseline"] = cpu_baseline(6, 10000, 1, margin=MARGIN["arm"])
    res["plan"] = plan_metric(ctx, B, 6, cpu=cpu)
    return res


def small_batch_metric(ctx):
    """The reference's own training batch (2 environments x 10 000 pairs per step, M:996,1060-1080) through
    Model.train()'s loop (pntf/trainer.py: GPU-resident table, fused gather): wall time per step over whole epochs."""
    import torch
    from pntf import synth
    from pntf.model import Model_multi
    from pntf.trainer import run_training
    d, rows = 3, 60000
    datas, Bs = [], []
    for e in range(2):
        c, s = synth.sample_pairs(rows, d, ctx.dev, seed=60 + e, margin=MARGIN["test"], near_obstacle=True)
        datas.append(torch.cat([c, s], dim=1))
        Bs.append(synth.sample_B(d, ctx.dev, seed=70 + e, E=1)[0])
    torch.manual_seed(0)
    model = Model_multi("", "", d, 2, device=str(ctx.dev))
    model.init_network()
    model.network.precision = ctx.prec
    run_training(model, datas, Bs, epochs=3, log=lambda s: None)                 # warm-up (allocations, first launches)
    torch.cuda.synchronize()
    epochs = 10
    t0 = time.perf_counter()
    hist = run_training(model, datas, Bs, epochs=epochs, log=lambda s: None)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    steps = epochs * 6
    return {"workload": "Model.train() loop at the reference's batch: 2 envs x 10 000 pairs per step, 6 steps per epoch",
            "ms_per_step_wall": dt * 1e3 / steps, "value": 20000 * steps / dt, "unit": UNIT, "epochs": epochs,
            "loss_first": hist[0][1], "loss_last": hist[-1][1]}


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
    # arm branch (:153-218): forward kinematics of a synthetic six-joint chain, 6 box links x 26 vertices per
    # configuration against the same mesh, minimum over the arm
    from pntf import arm_sampling as AS
    J = [("l%d" % i, (0.0, 0.02 * (i % 2), 0.12), (0.0, 1.5708 * (i % 2), 0.0), ((0, 0, 1), (0, 1, 0))[i % 2]) for i in range(6)]
    chain = AS.SerialChain([{"name": n_, "type": "revolute", "xyz": x_, "rpy": r_, "axis": a_} for n_, x_, r_, a_ in J], device=dev)
    verts = {n_: AS.box_link_vertices((-0.02, -0.02, 0.0), (0.02, 0.02, 0.12)) for n_, _, _, _ in J}
    th = (torch.rand(1 << 16, 6, device=dev, generator=g) - 0.5) * 6.2831855
    sec_arm = timed(lambda: AS.arm_obstacle_distance(th, chain, verts, mesh), reps=2)
    res["arm_configurations_per_s"] = th.shape[0] / sec_arm
    res["arm_vertex_queries_per_s"] = th.shape[0] * 5 * verts["l0"].shape[0] / sec_arm      # base and first link are skipped (:168)
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


def plan_metric(ctx, B, d, Q=65536, cpu=True):
    """Batched bidirectional path integration, paths/s (BASELINE configs[4]): 65 536 start/goal queries ~ U[-0.4,0.4]^d,
    reference step / threshold / cap (test/gib_plan.py:74-86, test/arm_plan.py:140-152), on freshly initialised weights.
    The queries are split evenly over the ranks (Model.plan(shard=True): contiguous slices, no collective); the time is
    the max over ranks, the rate the whole job's."""
    import torch
    from pntf import capi, synth
    from pntf.model import Model_multi
    from pntf.dist import shard_range
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    q_all = synth.sample_queries(Q, d, dev)
    lo, hi = shard_range(Q, rank, world)
    q = q_all[lo:hi].contiguous()
    eta, thr, cap = (0.03, 0.06, 500) if d == 3 else (0.015, 0.03, 300)
    torch.manual_seed(0)
    fresh = Model_multi("", "", d, 1, device=str(dev))
    fresh.init_network()
    flat = fresh.network.flat().flat
    Bq = B.reshape(-1, d, 128)[:1].contiguous()
    res = {"queries": Q, "queries_per_gpu": hi - lo, "n_gpus": world, "eta": eta, "thr": thr, "max_iter": cap}
    variants = [("compat", "tc", True), ("exact", "tc", True), ("exact", "tc", False), ("exact", "fp32_simt", True)] if d == 3 else \
        [("exact", "tc", True), ("exact", "tc", False), ("exact", "fp32_simt", True)]
    for mode, prec_name, fill in variants:
        if world > 1 and prec_name != "tc":
            continue
        fl = capi.GRAD_COMPAT if mode == "compat" else capi.GRAD_EXACT
        prec = capi.PREC_TC if prec_name == "tc" else capi.PREC_FP32
        if prec_name == "tc":      # warm-up at full size (the waypoint buffer comes from torch's caching allocator)
            capi.plan(flat, Bq, q, d, eta, thr, cap, fl, prec=prec, fill_tail=fill)
        barrier(ctx)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        way, nst = capi.plan(flat, Bq, q, d, eta, thr, cap, fl, prec=prec, fill_tail=fill)
        b.record()
        torch.cuda.synchronize()
        ms = max_over_ranks(ctx, a.elapsed_time(b))
        steps_t = nst.double().sum().reshape(1)
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(steps_t)
        steps = float(steps_t.item())
        live = (torch.arange(cap + 2, device=dev).view(1, -1) <= nst.view(-1, 1))
        finite = bool(torch.isfinite(way[live]).all().item())
        key = "%s_%s%s" % (mode, prec_name, "" if fill else "_compact")
        res[key] = {"paths_per_s": Q / (ms * 1e-3), "path_steps_per_s": steps / (ms * 1e-3), "mean_steps": steps / Q,
                    "ms": ms, "finite": finite, "tail_filled": fill,
                    "alg_waypoint_GBps": steps * 2 * d * 4 / (ms * 1e-3) / 1e9,
                    "written_GBps": (world * way.numel() * 4 if fill else steps * 2 * d * 4) / (ms * 1e-3) / 1e9,
                    "alg_tflops": steps * PLAN_FLOP_PER_STEP / (ms * 1e-3) / 1e12 if d == 3 else None}
        del way, nst
    if cpu and rank == 0:
        res["cpu_baseline"] = plan_cpu_baseline(q_all[:256].cpu(), Bq[0].cpu(), d, eta, thr, cap)
    return res


def plan_cpu_baseline(q, Bw, d, eta, thr, cap):
    """The reference plan loop on the host cores, 256 of the same queries, exact gradient, fp32: with oracle/_ref the
    reference's own Model.Gradient (d = 3: the row-wise batched M:1218-1248 with per-row termination masks, grad of
    out_grad -- exact flavour; d = 6: S:1247-1282 is batch-1 only, so 16 queries one at a time), else the oracle port."""
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
