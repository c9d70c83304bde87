#!/bin/bash
# One GPU pass: parity tests, smoke, bench (tensor-core path), reference arm, ncu launch list, ncu full capture.
# Usage: scripts/gpu_round.sh <tag> [skip-tests]
TAG=${1:-rXX}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu_$TAG.txt 2>&1
nproc >> gpurun_out/gpu_$TAG.txt
if [ -z "$2" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
  tail -5 gpurun_out/pytest_gpu_$TAG.log
  timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_$TAG.log
  tail -3 gpurun_out/smoke_$TAG.log
fi
timeout 900 python bench.py --prec tc --plan > gpurun_out/bench_tc_$TAG.json 2> gpurun_out/bench_tc_$TAG.err; echo "bench rc=$?"
cat gpurun_out/bench_tc_$TAG.json; tail -3 gpurun_out/bench_tc_$TAG.err
timeout 600 python bench.py --prec tc --d 6 --no-cpu-baseline > gpurun_out/bench_tc_d6_$TAG.json 2> gpurun_out/bench_tc_d6_$TAG.err
cat gpurun_out/bench_tc_d6_$TAG.json
CMD="python bench.py --prec tc --steps 1 --warmup 1 --pairs 262144 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_|simt_|pack_|adamw|loss_finalize|dw_kernel|FillFunctor" -c 400 --csv --log-file gpurun_out/launches_tc_$TAG.csv $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:tc_ -c 3 -o gpurun_out/tc_full_$TAG -f $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
