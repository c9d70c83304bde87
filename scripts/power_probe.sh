#!/bin/bash
# board vs module power under the training step: scripts/power_probe.sh  (prints nvidia-smi -q -d POWER samples taken while
# bench.py runs 60 timed steps)
cd "$(dirname "$0")/.."
python bench.py --no-cpu-baseline --no-e2e --no-plan --steps 60 --warmup 3 > /tmp/pp.json 2> /tmp/pp.err &
PID=$!
for i in $(seq 1 40); do
  sleep 0.5
  nvidia-smi --query-gpu=clocks.sm,power.draw,power.draw.instant,power.draw.average,module.power.draw.instant,module.power.draw.average,module.power.limit,clocks_event_reasons.sw_power_cap --format=csv,noheader 2>&1 | head -1
done
wait $PID
python -c "
import json
j=json.loads(open('/tmp/pp.json').read().strip().splitlines()[-1]); print(j['value'], j['ms_per_step'], j['clocks'])"
nvidia-smi -q -d POWER | sed -n 1,60p
