#!/bin/bash
# round 2, GPU call 8: L2 eviction policies (rows evict_first, records evict_last): quick parity tests, then A/B
mkdir -p gpurun_out
FMPNP_LIGHT=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_fullsize_parity.py -m gpu -x -q > gpurun_out/r02_c8_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c8_pytest.log
tail -3 gpurun_out/r02_c8_pytest.log
X="--no-e2e --no-cpu --sustain-s 0"
for P in 1 0; do for L in 1 0; do
  FMPNP_L2_POLICY=$P FMPNP_LIGHT=$L python bench.py $X > gpurun_out/r02_c8_p${P}_l${L}_cfg2.json 2>> gpurun_out/r02_c8.err
  FMPNP_L2_POLICY=$P FMPNP_LIGHT=$L python bench.py $X --no-every-row --config geman_lm > gpurun_out/r02_c8_p${P}_l${L}_cfg3.json 2>> gpurun_out/r02_c8.err
done; done
FMPNP_LIGHT=1 python bench.py $X --no-every-row --iters 100 > gpurun_out/r02_c8_p1_l1_it100.json 2>> gpurun_out/r02_c8.err
FMPNP_LIGHT=1 python bench.py $X --no-every-row --config pyramid > gpurun_out/r02_c8_p1_l1_cfg4.json 2>> gpurun_out/r02_c8.err
FMPNP_LIGHT=1 python bench.py $X --no-every-row --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c8_p1_l1_gs64rot.json 2>> gpurun_out/r02_c8.err
for f in gpurun_out/r02_c8_p*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry']['tasks_per_evaluation'], r.get('reused_fraction'), r.get('every_row_kernel_ms'))"; done
tail -5 gpurun_out/r02_c8.err
