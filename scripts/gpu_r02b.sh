#!/bin/bash
cd "$(dirname "$0")/.."
OUT=gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py -x -q -m gpu > $OUT/r02b_pytest_trainer.log 2>&1
echo "rc=$?" >> $OUT/r02b_pytest_trainer.log
tail -40 $OUT/r02b_pytest_trainer.log
timeout 1500 python scripts/explore_parity.py > $OUT/r02b_explore.jsonl 2> $OUT/r02b_explore.err
echo "explore rc=$?"; tail -3 $OUT/r02b_explore.err
