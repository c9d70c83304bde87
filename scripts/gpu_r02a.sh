#!/bin/bash
# round 2, run a: fused backward (reverse sweep + trailing weight gradients in one launch) -- parity, then A/B of the split
cd "$(dirname "$0")/.."
OUT=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $OUT/r02a_gpu.txt
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_fullsize.py -x -q -m gpu > $OUT/r02a_pytest.log 2>&1
echo "pytest rc=$?" >> $OUT/r02a_pytest.log
tail -5 $OUT/r02a_pytest.log
B="--no-cpu-baseline --no-plan --no-e2e --steps 5 --warmup 3"
for W in split 24 32 40 48; do
  if [ $W = split ]; then export PNTF_BWD_SPLIT=1; unset PNTF_DW_CTAS; else unset PNTF_BWD_SPLIT; export PNTF_DW_CTAS=$W; fi
  timeout 300 python bench.py $B > $OUT/r02a_bench_d3_$W.json 2> $OUT/r02a_bench_d3_$W.err
  python - <<PY
import json
try:
    j=json.loads(open("$OUT/r02a_bench_d3_$W.json").read().strip().splitlines()[-1])
    print("$W", "d3 Mpairs/s %.2f ms %.2f"%(j["value"]/1e6, j["ms_per_step"]), {k:round(v["ms_per_step"],2) for k,v in j["kernels"].items()}, j["clocks"]["sm_mhz"], j["final_loss_n"])
except Exception as e:
    print("$W failed", e)
PY
done
unset PNTF_BWD_SPLIT; export PNTF_DW_CTAS=32
timeout 300 python bench.py $B --d 6 > $OUT/r02a_bench_d6_32.json 2> $OUT/r02a_bench_d6_32.err
tail -c 600 $OUT/r02a_bench_d6_32.json
