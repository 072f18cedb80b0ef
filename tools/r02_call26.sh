#!/bin/bash
# round 2, GPU call 26: experiment -- persisting-L2 carve-out for the evict_last lines
mkdir -p gpurun_out
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for MB in 0 48 96 0 48 96; do
  FMPNP_PERSIST_L2_MB=$MB python bench.py $X >> gpurun_out/r02_c26_mb${MB}_cfg2.json 2>> gpurun_out/r02_c26.err
  FMPNP_PERSIST_L2_MB=$MB python bench.py $X --iters 100 >> gpurun_out/r02_c26_mb${MB}_it100.json 2>> gpurun_out/r02_c26.err
done
grep "persisting" gpurun_out/r02_c26.err | sort | uniq -c
for f in gpurun_out/r02_c26_*.json; do echo $f; python -c "
import json,sys
for l in open('$f').read().strip().splitlines():
    d=json.loads(l); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3))"; done
