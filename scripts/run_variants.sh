#!/bin/bash
# On the GPU box: for every variant library run a short bench and print the per-kernel times.
# Usage: scripts/run_variants.sh "<bench args>" tag1 tag2 ...
ARGS=$1; shift
mkdir -p gpurun_out
cp p-ntfields_b200/pntf/libpntf_b200.so /tmp/libpntf_orig.so
for T in "$@"; do
  cp p-ntfields_b200/pntf/variants/libpntf_b200.$T.so p-ntfields_b200/pntf/libpntf_b200.so
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-plan $ARGS > gpurun_out/var_$T.json 2> gpurun_out/var_$T.err
  python - "$T" <<'PY'
import json, sys
t = sys.argv[1]
try:
    j = json.loads(open("gpurun_out/var_%s.json" % t).read().strip().splitlines()[-1])
    k = j["kernels"]
    print("%-12s %.3f M pairs/s  step %.2f ms  fwd %.2f rev %.2f dw %.2f  loss %.6f" % (t, j["value"] / 1e6, j["ms_per_step"], k["tc_forward"]["ms_per_step"], k["tc_reverse"]["ms_per_step"], k["tc_dw"]["ms_per_step"], j["final_loss_n"]))
except Exception as e:
    print(t, "FAILED", e, open("gpurun_out/var_%s.err" % t).read()[-400:])
PY
done
cp /tmp/libpntf_orig.so p-ntfields_b200/pntf/libpntf_b200.so
