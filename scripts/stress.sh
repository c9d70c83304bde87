#!/bin/bash
# Reliability check: run a short bench N times, count failures.  Usage: scripts/stress.sh N "<bench args>"
N=${1:-20}; ARGS=${2:---prec tc --steps 1 --warmup 1 --pairs 262144 --no-cpu-baseline --no-e2e}
fail=0
for i in $(seq 1 $N); do
  s=$(date +%s.%N)
  timeout 300 python bench.py $ARGS > /tmp/stress_$i.json 2> /tmp/stress_$i.err
  rc=$?
  e=$(date +%s.%N)
  if [ $rc -ne 0 ]; then fail=$((fail+1)); echo "run $i rc=$rc $(python -c "print(round($e - $s, 1))") s: $(grep -n "bench.py\", line" /tmp/stress_$i.err | tail -2 | tr '\n' ' ' | cut -c1-200) $(tail -1 /tmp/stress_$i.err | cut -c1-120)"; fi
done
echo "stress: $fail failures of $N"
