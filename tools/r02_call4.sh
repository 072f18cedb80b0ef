#!/bin/bash
# round 2, GPU call 4: the light-evaluation route (fmpnp_light.cuh): tests, then cfg2 / cfg3 / cfg4 with FMPNP_LIGHT=0 / 1
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c4_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c4_pytest.log
tail -5 gpurun_out/r02_c4_pytest.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for L in 0 1; do
  FMPNP_LIGHT=$L python bench.py $X > gpurun_out/r02_c4_l${L}_cfg2.json 2>> gpurun_out/r02_c4.err
  FMPNP_LIGHT=$L python bench.py $X --config geman_lm > gpurun_out/r02_c4_l${L}_cfg3.json 2>> gpurun_out/r02_c4.err
  FMPNP_LIGHT=$L python bench.py $X --config pyramid > gpurun_out/r02_c4_l${L}_cfg4.json 2>> gpurun_out/r02_c4.err
  FMPNP_LIGHT=$L python bench.py $X --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c4_l${L}_gs64rot.json 2>> gpurun_out/r02_c4.err
  FMPNP_LIGHT=$L python bench.py $X --queries 32 > gpurun_out/r02_c4_l${L}_q32.json 2>> gpurun_out/r02_c4.err
done
for f in gpurun_out/r02_c4_l*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry'], r.get('reused_fraction'))"; done
tail -5 gpurun_out/r02_c4.err
