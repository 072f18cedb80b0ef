#!/bin/bash
# round 2, GPU call 21: experiment -- the owner loop with the general evaluation (stream_points + reduce_points) instead of light_evaluate
mkdir -p gpurun_out
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for L in 1 2 1 2; do
  FMPNP_LIGHT=$L python bench.py $X >> gpurun_out/r02_c21_l${L}_cfg2.json 2>> gpurun_out/r02_c21.err
done
FMPNP_LIGHT=2 python bench.py $X --config geman_lm > gpurun_out/r02_c21_l2_cfg3.json 2>> gpurun_out/r02_c21.err
FMPNP_LIGHT=1 python bench.py $X --config geman_lm > gpurun_out/r02_c21_l1_cfg3.json 2>> gpurun_out/r02_c21.err
python bench.py $X --config geman_lm > gpurun_out/r02_c21_auto_cfg3.json 2>> gpurun_out/r02_c21.err
for f in gpurun_out/r02_c21_*.json; do echo $f; python -c "
import json,sys
for l in open('$f').read().strip().splitlines():
    d=json.loads(l); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3))"; done
