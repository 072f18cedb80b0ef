#!/bin/bash
# round 2, GPU call 3: occupancy experiment -- the task kernel compiled for 2 / 3 / 4 CTAs per SM (128 / 80 / 64 registers)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py -m gpu -x -q > gpurun_out/r02_c3_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c3_pytest.log
FMPNP_CTAS_PER_SM=4 python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py -m gpu -x -q > gpurun_out/r02_c3_pytest_ps4.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c3_pytest_ps4.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for PS in 2 3 4; do
  for IT in 0 20; do
    FMPNP_CTAS_PER_SM=$PS python bench.py $X --iters $IT > gpurun_out/r02_c3_ps${PS}_it${IT}.json 2>> gpurun_out/r02_c3.err
  done
done
for PS in 3 4; do
  FMPNP_CTAS_PER_SM=$PS FMPNP_UNITS=4 python bench.py $X --iters 20 > gpurun_out/r02_c3_ps${PS}_u4_it20.json 2>> gpurun_out/r02_c3.err
  FMPNP_CTAS_PER_SM=$PS python bench.py $X --config geman_lm > gpurun_out/r02_c3_ps${PS}_cfg3.json 2>> gpurun_out/r02_c3.err
done
tail -3 gpurun_out/r02_c3_pytest.log gpurun_out/r02_c3_pytest_ps4.log
for f in gpurun_out/r02_c3_ps*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry'])"; done
tail -5 gpurun_out/r02_c3.err
