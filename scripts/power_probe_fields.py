"""Board power and SM clock while ONE inference kernel runs back to back (no stash traffic): python scripts/power_probe_fields.py
Prints, per kernel, the per-launch time and the nvidia-smi samples (clocks.sm, power.draw.instant, sw_power_cap) taken meanwhile."""
import os, subprocess, sys, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import capi, synth
from pntf.model import Model_multi
dev = torch.device("cuda:0")
d, n = 3, 1 << 20
torch.manual_seed(0)
m = Model_multi("", "", d, 1, device=str(dev)); m.init_network()
flat = m.network.flat().flat
B = synth.sample_B(d, dev, seed=1, E=1).contiguous()
coords, _ = synth.sample_pairs(n, d, dev, seed=2, margin=0.5 / 12)
for name, kind, reps in (("TravelTimes", capi.FIELD_TRAVEL_TIME, 1200), ("Speed", capi.FIELD_SPEED, 400), ("Speed2", capi.FIELD_SPEED2, 300)):
    capi.field(kind, flat, B, coords, 1, n, d, prec=capi.PREC_TC)
    torch.cuda.synchronize()
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw.instant,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits",
                          "-lms", "250"], stdout=subprocess.PIPE, text=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        capi.field(kind, flat, B, coords, 1, n, d, prec=capi.PREC_TC)
    b.record(); torch.cuda.synchronize()
    p.terminate()
    out = p.communicate()[0].strip().splitlines()
    print("%s: %.3f ms per 1 Mi pairs over %.1f s; samples (MHz, W, sw_power_cap): %s" % (name, a.elapsed_time(b) / reps, a.elapsed_time(b) / 1e3,
                                                                                            " | ".join(out[2:-1][:14])), flush=True)
    time.sleep(2)
