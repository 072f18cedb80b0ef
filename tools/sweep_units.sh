#!/bin/bash
# Small-batch crossover (scheduler vs clusters) and tasks-per-evaluation sweep.  Usage: tools/sweep_units.sh <tag>
tag=${1:-units}
out=gpurun_out/${tag}.txt
: > $out
one() {
  echo "== $*" >> $out
  env "$1" "$2" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e "${@:3}" 2>>gpurun_out/${tag}.err | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'])
" >> $out
}
for b in 2 4 8 16; do
  for sch in 1 0; do one FMPNP_SCHEDULER=$sch X=0 --config geman_lm --queries $b; done
done
for u in 1 2 4 8 16; do
  one FMPNP_UNITS=$u X=0 --queries 256
  one FMPNP_UNITS=$u X=0 --config pyramid --points 8192 --queries 256
  one FMPNP_UNITS=$u X=0 --config geman_lm --queries 64
done
for ps in 1 2; do one FMPNP_CTAS_PER_SM=$ps X=0 --config geman_lm --queries 16; done
cat $out
