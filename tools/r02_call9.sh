#!/bin/bash
# round 2, GPU call 9: pow special cases, 128-bit tree reads: parity subset, cfg2 / cfg3 with the light route on / off
mkdir -p gpurun_out
FMPNP_LIGHT=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_fullsize_parity.py -m gpu -x -q > gpurun_out/r02_c9_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c9_pytest.log
tail -3 gpurun_out/r02_c9_pytest.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for L in 1 0; do
  FMPNP_LIGHT=$L python bench.py $X > gpurun_out/r02_c9_l${L}_cfg2.json 2>> gpurun_out/r02_c9.err
  FMPNP_LIGHT=$L python bench.py $X --config geman_lm > gpurun_out/r02_c9_l${L}_cfg3.json 2>> gpurun_out/r02_c9.err
  FMPNP_LIGHT=$L python bench.py $X --config pyramid > gpurun_out/r02_c9_l${L}_cfg4.json 2>> gpurun_out/r02_c9.err
done
for f in gpurun_out/r02_c9_l*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry']['tasks_per_evaluation'], r.get('reused_fraction'))"; done
tail -5 gpurun_out/r02_c9.err
