#!/bin/bash
cd "$(dirname "$0")/.."
OUT=gpurun_out
nvidia-smi -L | head -4
timeout 300 python scripts/debug_small_batch.py > $OUT/r02f_small.log 2>&1; tail -14 $OUT/r02f_small.log
timeout 900 python -m pytest tests/test_gpu_dist.py tests/test_gpu_north_star.py -x -q -m gpu > $OUT/r02f_pytest.log 2>&1
echo "rc=$?" >> $OUT/r02f_pytest.log; tail -25 $OUT/r02f_pytest.log
for N in 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > $OUT/r02f_bench_${N}gpu.json 2> $OUT/r02f_bench_${N}gpu.err
echo "bench $N rc=$?"; tail -2 $OUT/r02f_bench_${N}gpu.err; head -c 600 $OUT/r02f_bench_${N}gpu.json
done
