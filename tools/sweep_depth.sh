#!/bin/bash
# Ring depth / chunk size sweep.  Usage: tools/sweep_depth.sh <tag>
tag=${1:-depth}
out=gpurun_out/${tag}.txt
: > $out
one() {
  echo "== $*" >> $out
  env "$1" "$2" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e "${@:3}" 2>>gpurun_out/${tag}.err | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['kernel_ms'])
" >> $out
}
for d in 2 3 4; do
  one FMPNP_MIN_DEPTH=$d X=0 --queries 256
  one FMPNP_MIN_DEPTH=$d X=0 --config pyramid --points 8192 --queries 256
done
one FMPNP_MIN_DEPTH=3 FMPNP_RING_KB=100 --queries 256
one FMPNP_CTAS_PER_SM=1 X=0 --queries 256
one FMPNP_CTAS_PER_SM=1 FMPNP_MIN_DEPTH=4 --queries 256
cat $out
