#!/bin/bash
# A/B of two builds of the library on the same box.  Usage: tools/sweep_ab.sh <tag> [prev.so]
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
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'])
" >> $out
  done
  unset FMPNP_LIB
}
one --config geman_lm --channels 1024 --points 256 --map 64 --queries 2048
one --config geman_lm --channels 2048 --points 512 --map 64 --queries 1024
one --config geman_lm --channels 2048 --points 128 --map 64 --queries 1024
one --queries 256
cat $out
