#!/bin/bash
# round 2: trainer tests, smoke, ncu launch list + full capture of the training kernels and of the inference forward
cd "$(dirname "$0")/.."
OUT=gpurun_out
timeout 900 python -m pytest tests/test_gpu_trainer.py -x -q -m gpu 2>&1 | tail -4
timeout 300 python __graft_entry__.py --smoke > $OUT/r02j_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/r02j_smoke.log
CMD="python bench.py --prec tc --steps 1 --warmup 1 --pairs 262144 --no-cpu-baseline --no-e2e --no-plan"
timeout 300 $CMD > $OUT/r02j_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_|simt_|pack_|adamw|loss_finalize|dw_kernel" -c 400 --csv --log-file $OUT/r02j_launches_tc_262144pairs.csv $CMD > $OUT/r02j_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:tc_ -c 3 -o $OUT/r02j_tc_full -f $CMD > $OUT/r02j_ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 300 python scripts/fields_once.py > $OUT/r02j_fields_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 2 -o $OUT/r02j_fields_full -f python scripts/fields_once.py > $OUT/r02j_ncu_fields.log 2>&1
echo "ncu fields rc=$?"
ls -la $OUT/*.ncu-rep
