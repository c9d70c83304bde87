#!/bin/bash
# quick TC iteration: TC parity tests + short bench.  Usage: scripts/gpu_tc.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -x -q -k "tc" > gpurun_out/pytest_tc_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_tc_$TAG.log
tail -15 gpurun_out/pytest_tc_$TAG.log
timeout 300 python bench.py --prec tc --no-cpu-baseline --no-e2e > gpurun_out/bench_tcq_$TAG.json 2> gpurun_out/bench_tcq_$TAG.err; echo "bench rc=$?"
python - <<PY
import json
try:
    j=json.load(open("gpurun_out/bench_tcq_$TAG.json"))
    print("value %.3e ms/step %.2f whole-step TF %.1f"%(j["value"], j["ms_per_step"], j["roofline"]["whole_step_tflops"]))
    print({k:round(v["ms_per_step"],2) for k,v in j["kernels"].items()}, "loss", j["final_loss_n"])
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_tcq_$TAG.err").read()[-2000:])
PY
