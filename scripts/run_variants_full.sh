#!/bin/bash
# On the GPU box: for every variant library run the default bench (training step + plan + fields) and print one line.
# Usage: scripts/run_variants_full.sh "<bench args>" tag1 tag2 ...
ARGS=$1; shift
mkdir -p gpurun_out
cp p-ntfields_b200/pntf/libpntf_b200.so /tmp/libpntf_orig.so
for T in "$@"; do
  cp p-ntfields_b200/pntf/variants/libpntf_b200.$T.so p-ntfields_b200/pntf/libpntf_b200.so
  timeout 600 python bench.py --no-cpu-baseline --no-e2e $ARGS > gpurun_out/varf_$T.json 2> gpurun_out/varf_$T.err
  python - "$T" <<'PY'
import json, sys
t = sys.argv[1]
try:
    j = json.loads(open("gpurun_out/varf_%s.json" % t).read().strip().splitlines()[-1])
    k = j["kernels"]
    if "extra" not in j:      # --no-plan: training step only
        print("%-8s %.3f M pairs/s fwd %.2f rev %.2f dw %.2f | loss %.6f | %d MHz" % (
            t, j["value"] / 1e6, k["tc_forward"]["ms_per_step"], k["tc_reverse"]["ms_per_step"], k["tc_dw"]["ms_per_step"],
            j["final_loss_n"], j["clocks"]["sm_mhz"]))
        raise SystemExit
    f = j["extra"]["fields"]; p = j["extra"]["plan"]["exact_tc"]
    print("%-8s %.3f M pairs/s fwd %.2f rev %.2f dw %.2f | plan %.3f M paths/s | TT %.1f speed %.1f speed2 %.1f lap %.1f M/s | loss %.6f | %d MHz" % (
        t, j["value"] / 1e6, k["tc_forward"]["ms_per_step"], k["tc_reverse"]["ms_per_step"], k["tc_dw"]["ms_per_step"],
        p["paths_per_s"] / 1e6, f["travel_times_pairs_per_s"] / 1e6, f["speed_pairs_per_s"] / 1e6, f["speed2_pairs_per_s"] / 1e6,
        f["out_laplace_pairs_per_s"] / 1e6, j["final_loss_n"], j["clocks"]["sm_mhz"]))
except Exception as e:
    print(t, "FAILED", e, open("gpurun_out/varf_%s.err" % t).read()[-400:])
PY
done
cp /tmp/libpntf_orig.so p-ntfields_b200/pntf/libpntf_b200.so
