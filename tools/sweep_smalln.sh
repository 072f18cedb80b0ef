#!/bin/bash
# Many queries with few points each (what RANSAC inlier sets look like).  Usage: tools/sweep_smalln.sh <tag>
tag=${1:-smalln}
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
one --config pyramid --channels 2048 --points 512 --map 64 --queries 1024
one --config geman_lm --channels 2048 --points 512 --map 64 --queries 1024
one --config geman_lm --channels 2048 --points 128 --map 64 --queries 1024
one --config geman_lm --channels 1024 --points 256 --map 64 --queries 2048
one --config geman_lm --channels 2048 --points 512 --map 64 --queries 64
cat $out
