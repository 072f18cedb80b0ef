#!/bin/bash
# Small-batch latency: scheduler vs clusters, 1 vs 2 CTAs per SM.  Usage: tools/sweep_small.sh <tag>
tag=${1:-small}
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
for b in 1 2 4 8 16 24 32 48; do
  one FMPNP_SCHEDULER=1 FMPNP_CTAS_PER_SM=1 --config geman_lm --queries $b
  one FMPNP_SCHEDULER=1 FMPNP_CTAS_PER_SM=2 --config geman_lm --queries $b
  [ $b -le 16 ] && one FMPNP_SCHEDULER=0 X=0 --config geman_lm --queries $b
done
one X=0 Y=0 --queries 256
one X=0 Y=0 --config pyramid --points 8192 --queries 256
cat $out
