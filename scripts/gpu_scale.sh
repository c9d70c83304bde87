#!/bin/bash
# data-parallel bench at N ranks (weak + strong scaling, sharded plan): scripts/gpu_scale.sh <tag> <N...>
cd "$(dirname "$0")/.."
OUT=gpurun_out; TAG=$1; shift
for N in "$@"; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_${N}gpu.json 2> $OUT/${TAG}_bench_${N}gpu.err
  echo "bench $N rc=$?"
  python - <<PY
import json
try:
    j=json.loads(open("$OUT/${TAG}_bench_${N}gpu.json").read().strip().splitlines()[-1])
    ex=j["extra"]
    print($N, "weak %.2f M pairs/s %.2f ms"%(j["value"]/1e6,j["ms_per_step"]), "e2e %.2f"%(j["e2e"]["value"]/1e6), "strong %.2f M %.2f ms"%(ex["strong"]["value"]/1e6, ex["strong"]["ms_per_step"]),
          "plan", {k:round(v["paths_per_s"]/1e6,2) for k,v in ex["plan"].items() if isinstance(v,dict) and "ms" in v}, j["clocks"])
except Exception as e:
    print("failed", e)
PY
done
