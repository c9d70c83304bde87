#!/bin/bash
cd "$(dirname "$0")/.."
OUT=gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > $OUT/r02d_pytest.log 2>&1
echo "rc=$?" >> $OUT/r02d_pytest.log
tail -30 $OUT/r02d_pytest.log
timeout 300 python scripts/debug_train_nan.py > $OUT/r02d_nan.log 2>&1
grep -c "^tc" $OUT/r02d_nan.log; tail -2 $OUT/r02d_nan.log | cut -c1-900
