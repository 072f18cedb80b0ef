#!/bin/bash
# round 2, GPU call 29: tasks per evaluation on the task route after the round's changes (cfg3, cfg4), light-route miss threshold on a moving workload
mkdir -p gpurun_out
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for U in 1 2 4; do
  FMPNP_UNITS=$U python bench.py $X --config geman_lm > gpurun_out/r02_c29_u${U}_cfg3.json 2>> gpurun_out/r02_c29.err
  FMPNP_UNITS=$U python bench.py $X --config pyramid > gpurun_out/r02_c29_u${U}_cfg4.json 2>> gpurun_out/r02_c29.err
done
for D in 32 128; do
  FMPNP_LIGHT_MISS_DEN=$D python bench.py $X --grad-scale 0.125 > gpurun_out/r02_c29_den${D}_gs8.json 2>> gpurun_out/r02_c29.err
  FMPNP_LIGHT_MISS_DEN=$D python bench.py $X > gpurun_out/r02_c29_den${D}_cfg2.json 2>> gpurun_out/r02_c29.err
done
for f in gpurun_out/r02_c29_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry']['tasks_per_evaluation'])"; done
