#!/bin/bash
# round 2, GPU call 20: light tasks (light_project inside the tasks of the task route): parity subset, A/B on cfg2 / cfg3 / cfg4 and moving workloads
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_edge_and_scale.py tests/test_gpu_fullsize_parity.py -m gpu -x -q > gpurun_out/r02_c20_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c20_pytest.log
tail -n 3 gpurun_out/r02_c20_pytest.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for LT in 1 0; do
  export FMPNP_LIGHT_TASKS=$LT
  python bench.py $X > gpurun_out/r02_c20_lt${LT}_cfg2.json 2>> gpurun_out/r02_c20.err
  python bench.py $X --config geman_lm > gpurun_out/r02_c20_lt${LT}_cfg3.json 2>> gpurun_out/r02_c20.err
  python bench.py $X --config pyramid > gpurun_out/r02_c20_lt${LT}_cfg4.json 2>> gpurun_out/r02_c20.err
  python bench.py $X --config geman_lm --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c20_lt${LT}_cfg3_gs64rot.json 2>> gpurun_out/r02_c20.err
  python bench.py $X --grad-scale 0.125 > gpurun_out/r02_c20_lt${LT}_gs8.json 2>> gpurun_out/r02_c20.err
  FMPNP_LIGHT=0 python bench.py $X > gpurun_out/r02_c20_lt${LT}_cfg2_l0.json 2>> gpurun_out/r02_c20.err
done
for f in gpurun_out/r02_c20_lt*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r.get('reused_fraction'))"; done
tail -n 3 gpurun_out/r02_c20.err
