#!/bin/bash
# r02w (was r02s): evidence pass on the round's final code: all GPU tests, smoke, default bench, ncu launch list and full captures
# (training kernels on one 12 478-tile chunk; the no-stash forward of the inference fields).
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
bash scripts/gpu_full.sh r02w 2>&1 | tee $OUT/r02w_full.txt
timeout 300 python __graft_entry__.py --smoke > $OUT/r02w_smoke.log 2>&1; echo "smoke rc=$?"
CMD="python bench.py --prec tc --steps 1 --warmup 1 --pairs 149736 --no-cpu-baseline --no-e2e --no-plan"
timeout 300 $CMD > $OUT/r02w_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_|simt_|pack_|adamw|loss_finalize|dw_kernel" -c 400 --csv --log-file $OUT/r02w_launches_tc_149736pairs.csv $CMD > $OUT/r02w_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:tc_ -c 3 -o $OUT/r02w_tc_full -f $CMD > $OUT/r02w_ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 300 python scripts/fields_once.py > $OUT/r02w_fields_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 2 -o $OUT/r02w_fields_full -f python scripts/fields_once.py > $OUT/r02w_ncu_fields.log 2>&1
echo "ncu fields rc=$?"
ls -la $OUT/*.ncu-rep
bash scripts/stress.sh 12 "--prec tc --steps 1 --warmup 1 --pairs 262144 --no-cpu-baseline --no-e2e --no-plan" 2>&1 | tail -3 | tee gpurun_out/r02w_stress.txt
