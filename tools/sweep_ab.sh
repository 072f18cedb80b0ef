#!/bin/bash
# A/B of two builds of the library on the same box.  Usage: tools/sweep_ab.sh <tag> [prev.so] -- bench args are the cases below
tag=${1:-ab}
prev=${2:-featuremetric-pnp_b200/lib/libfmpnp_prev.so}
out=gpurun_out/${tag}.txt
: > $out
one() {
  for lib in new prev; do
    if [ $lib = prev ]; then export FMPNP_LIB=$PWD/$prev; else unset FMPNP_LIB; fi
    echo "== $lib $*" >> $out
    timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e "$@" 2>>gpurun_out/${tag}.err | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['kernel_ms'], d['clocks'])
" >> $out
  done
  unset FMPNP_LIB
}
one --config geman_lm --queries 256
one --config geman_lm --points 8192 --channels 128 --queries 296
cat $out
