#!/bin/bash
# full GPU validation: every -m gpu test, smoke, default bench.  Usage: scripts/gpu_full.sh <tag>
cd "$(dirname "$0")/.."
OUT=gpurun_out; TAG=${1:-rXX}
timeout 1800 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest.log 2>&1
echo "rc=$?" >> $OUT/${TAG}_pytest.log
tail -12 $OUT/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"; tail -2 $OUT/${TAG}_bench.err
python - <<PY
import json
j=json.loads(open("$OUT/${TAG}_bench.json").read().strip().splitlines()[-1])
print(j["value"], j["ms_per_step"], j["clocks"], {k:round(v["ms_per_step"],2) for k,v in j["kernels"].items()})
print("roof", j["roofline"]["frac"], j["roofline"]["traffic"], j["roofline"]["traffic_source"], "cpu", j["cpu_baseline"]["value"], j["cpu_baseline"]["kind"])
print({k:round(v/1e6,1) for k,v in j["extra"]["fields"].items() if k!="pairs"})
print("plan", {k:round(v["paths_per_s"]/1e6,2) for k,v in j["extra"]["plan"].items() if isinstance(v,dict) and "ms" in v})
print("d6", j["extra"]["d6"]["value"], "small", j["extra"]["small_batch"]["ms_per_step_wall"])
PY
