#!/bin/bash
# evidence pass (tags r02s, r02w, r02zf): scripts/gpu_r02w.sh <tag>; on the round's final code: all GPU tests, smoke, default bench, ncu launch list and full captures
# (training kernels on one 12 478-tile chunk; the no-stash forward of the inference fields).
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02w}
bash scripts/gpu_full.sh $TAG 2>&1 | tee $OUT/${TAG}_full.txt
timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
CMD="python bench.py --prec tc --steps 1 --warmup 1 --pairs 149736 --no-cpu-baseline --no-e2e --no-plan"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_|simt_|pack_|adamw|loss_finalize|dw_kernel" -c 400 --csv --log-file $OUT/${TAG}_launches_tc_149736pairs.csv $CMD > $OUT/${TAG}_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:tc_ -c 3 -o $OUT/${TAG}_tc_full -f $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 300 python scripts/fields_once.py > $OUT/${TAG}_fields_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_forward -c 2 -o $OUT/${TAG}_fields_full -f python scripts/fields_once.py > $OUT/${TAG}_ncu_fields.log 2>&1
echo "ncu fields rc=$?"
ls -la $OUT/*.ncu-rep
bash scripts/stress.sh 12 "--prec tc --steps 1 --warmup 1 --pairs 262144 --no-cpu-baseline --no-e2e --no-plan" 2>&1 | tail -3 | tee gpurun_out/${TAG}_stress.txt
