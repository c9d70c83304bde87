#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/debug_tc.py > gpurun_out/debug_tc.log 2>&1; echo "debug rc=$?" >> gpurun_out/debug_tc.log
cat gpurun_out/debug_tc.log
timeout 900 python -m pytest tests/test_gpu_tc.py -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_tc.log
tail -15 gpurun_out/pytest_tc.log
