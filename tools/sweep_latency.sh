#!/bin/bash
# Small-batch latency: scheduler (auto) vs clusters, with and without a ratio test.  Usage: tools/sweep_latency.sh <tag>
tag=${1:-lat}
out=gpurun_out/${tag}.txt
: > $out
one() {
  echo "== $*" >> $out
  env "$1" "$2" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e "${@:3}" 2>>gpurun_out/${tag}.err | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['geometry'])
" >> $out
}
for b in 1 2 4 8 16; do
  one X=0 Y=0 --config geman_lm --queries $b
  one X=0 Y=0 --queries $b
  one FMPNP_SCHEDULER=0 X=0 --queries $b
done
one X=0 Y=0 --queries 32
one X=0 Y=0 --queries 256
one X=0 Y=0 --config geman_lm --queries 256
one X=0 Y=0 --config pyramid --points 8192 --queries 256
cat $out
