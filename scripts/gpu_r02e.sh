#!/bin/bash
cd "$(dirname "$0")/.."
OUT=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/r02e_pytest.log 2>&1
echo "rc=$?" >> $OUT/r02e_pytest.log
tail -30 $OUT/r02e_pytest.log
timeout 600 python bench.py > $OUT/r02e_bench_default.json 2> $OUT/r02e_bench_default.err
echo "bench rc=$?"; tail -3 $OUT/r02e_bench_default.err; head -c 1500 $OUT/r02e_bench_default.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/r02e_bench_reference.json 2> $OUT/r02e_bench_reference.err
echo "ref rc=$?"; cat $OUT/r02e_bench_reference.json | head -c 900
