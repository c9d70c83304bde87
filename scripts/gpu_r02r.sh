#!/bin/bash
# r02r: operand regeneration in the weight-gradient kernel (X_0, X_6 not stashed): full validation, same-box A/B against
# -DPNTF_STASH_REGEN=0, ncu launch list and full capture of the three training kernels on one 12 478-tile chunk.
cd "$(dirname "$0")/.."
OUT=gpurun_out; mkdir -p $OUT
bash scripts/gpu_full.sh r02r 2>&1 | tee $OUT/r02r_full.txt
bash scripts/run_variants_full.sh "" noregen regen noregen regen 2>&1 | tee $OUT/r02r_variants.txt
CMD="python bench.py --prec tc --steps 1 --warmup 1 --pairs 149736 --no-cpu-baseline --no-e2e --no-plan"
timeout 300 $CMD > $OUT/r02r_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_|simt_|pack_|adamw|loss_finalize|dw_kernel" -c 400 --csv --log-file $OUT/r02r_launches_tc_149736pairs.csv $CMD > $OUT/r02r_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:tc_ -c 3 -o $OUT/r02r_tc_full -f $CMD > $OUT/r02r_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT/*.ncu-rep
