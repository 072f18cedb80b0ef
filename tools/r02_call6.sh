#!/bin/bash
# round 2, GPU call 6: light route with staged loads: tests (route forced on for every batch size, then default), benches
mkdir -p gpurun_out
FMPNP_LIGHT=1 python -m pytest tests -m gpu -x -q > gpurun_out/r02_c6_pytest_light1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c6_pytest_light1.log
tail -4 gpurun_out/r02_c6_pytest_light1.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for L in 1 0; do
  FMPNP_LIGHT=$L python bench.py $X > gpurun_out/r02_c6_l${L}_cfg2.json 2>> gpurun_out/r02_c6.err
  FMPNP_LIGHT=$L python bench.py $X --config geman_lm > gpurun_out/r02_c6_l${L}_cfg3.json 2>> gpurun_out/r02_c6.err
  FMPNP_LIGHT=$L python bench.py $X --config pyramid > gpurun_out/r02_c6_l${L}_cfg4.json 2>> gpurun_out/r02_c6.err
  FMPNP_LIGHT=$L python bench.py $X --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c6_l${L}_gs64rot.json 2>> gpurun_out/r02_c6.err
  FMPNP_LIGHT=$L python bench.py $X --queries 32 > gpurun_out/r02_c6_l${L}_q32.json 2>> gpurun_out/r02_c6.err
done
FMPNP_LIGHT=1 python bench.py $X --iters 0 > gpurun_out/r02_c6_l1_it0.json 2>> gpurun_out/r02_c6.err
for U in 4 8; do FMPNP_LIGHT=1 FMPNP_UNITS=$U python bench.py $X > gpurun_out/r02_c6_l1_u${U}_cfg2.json 2>> gpurun_out/r02_c6.err; done
for f in gpurun_out/r02_c6_l*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry'], r.get('reused_fraction'))"; done
tail -5 gpurun_out/r02_c6.err
