#!/bin/bash
# Scheduler vs per-query geometry sweep (device-resident value only).  Usage: tools/sweep_sched.sh <tag>
tag=${1:-sweep}
out=gpurun_out/${tag}.txt
: > $out
run() {
  for sch in 1 0; do
    echo "== FMPNP_SCHEDULER=$sch $*" >> $out
    FMPNP_SCHEDULER=$sch timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e "$@" 2>>gpurun_out/${tag}.err | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['config'].get('queries'))
" >> $out
  done
}
run --queries 256
run --queries 296
run --queries 300
run --config geman_lm --queries 256
run --config geman_lm --queries 32
run --config geman_lm --queries 1
run --config pyramid --points 8192 --queries 256
run --config pyramid --points 8192 --queries 296
cat $out
