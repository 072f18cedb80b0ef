#!/bin/bash
# The reference's own pyramid shapes (input_configs/*.gin) at full size.  Usage: tools/sweep_real.sh <tag>
tag=${1:-real}
out=gpurun_out/${tag}.txt
: > $out
one() {
  echo "== $*" >> $out
  timeout 400 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e "$@" 2>>gpurun_out/${tag}.err | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['geometry'])
" >> $out
}
one --config pyramid --channels 2048 --queries 148
one --config geman_lm --channels 2048 --queries 148
one --config pyramid --channels 2048 --points 8192 --queries 100
cat $out
