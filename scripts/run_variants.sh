#!/bin/bash
# On the GPU box: for every variant library (scripts/build_variants.py) run the training-step bench and print one line.
# Usage: scripts/run_variants.sh "<extra bench args>" tag1 tag2 ...   (tags may repeat: same-box A/B/A/B)
ARGS=$1; shift
mkdir -p gpurun_out
cp p-ntfields_b200/pntf/libpntf_b200.so /tmp/libpntf_orig.so
n=0
for T in "$@"; do
  n=$((n+1))
  cp p-ntfields_b200/pntf/variants/libpntf_b200.$T.so p-ntfields_b200/pntf/libpntf_b200.so
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-plan $ARGS > gpurun_out/var_${T}_$n.json 2> gpurun_out/var_${T}_$n.err
  python - "$T" "$n" <<'PY'
import json, sys
t, n = sys.argv[1], sys.argv[2]
try:
    j = json.loads(open("gpurun_out/var_%s_%s.json" % (t, n)).read().strip().splitlines()[-1])
    k = j["kernels"]
    print("%-10s %.3f M pairs/s %.2f ms | %s | loss %.6f | %s MHz" % (t, j["value"] / 1e6, j["ms_per_step"],
          " ".join("%s %.2f" % (a.replace("tc_", ""), b["ms_per_step"]) for a, b in k.items() if a.startswith("tc_")), j["final_loss_n"], j["clocks"]["sm_mhz"]))
except Exception as e:
    print(t, "FAILED", e, open("gpurun_out/var_%s_%s.err" % (t, n)).read()[-400:])
PY
done
cp /tmp/libpntf_orig.so p-ntfields_b200/pntf/libpntf_b200.so
