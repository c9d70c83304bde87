#!/bin/bash
cd "$(dirname "$0")/.."
OUT=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/r02i_pytest.log 2>&1
echo "rc=$?" >> $OUT/r02i_pytest.log
tail -25 $OUT/r02i_pytest.log
timeout 600 python bench.py --no-cpu-baseline > $OUT/r02i_bench.json 2> $OUT/r02i_bench.err
echo "bench rc=$?"; tail -2 $OUT/r02i_bench.err
python - <<PY
import json
j=json.loads(open("$OUT/r02i_bench.json").read().strip().splitlines()[-1])
print(j["value"], j["ms_per_step"], j["clocks"], {k:round(v["ms_per_step"],2) for k,v in j["kernels"].items()})
print(j["extra"]["small_batch"])
PY
PNTF_TRAIN_GRAPH=0 timeout 300 python - <<PY
import sys,json
sys.argv=["bench.py"]
sys.path.insert(0,".")
import bench, torch
ctx=bench.Ctx(); ctx.world=1; ctx.rank=0; ctx.dev=torch.device("cuda:0"); ctx.prec="tc"
print("eager small_batch", bench.small_batch_metric(ctx))
PY
