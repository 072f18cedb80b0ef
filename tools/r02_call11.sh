#!/bin/bash
# round 2, GPU call 11: full GPU suite (default policy), parity subset with the light route forced, cfg2 / cfg3 / cfg4 lines
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c11_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c11_pytest.log
tail -3 gpurun_out/r02_c11_pytest.log
FMPNP_LIGHT=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_edge_and_scale.py -m gpu -x -q > gpurun_out/r02_c11_pytest_light1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c11_pytest_light1.log
tail -3 gpurun_out/r02_c11_pytest_light1.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
python bench.py $X > gpurun_out/r02_c11_cfg2.json 2>> gpurun_out/r02_c11.err
python bench.py $X --config geman_lm > gpurun_out/r02_c11_cfg3.json 2>> gpurun_out/r02_c11.err
python bench.py $X --config pyramid > gpurun_out/r02_c11_cfg4.json 2>> gpurun_out/r02_c11.err
FMPNP_LIGHT=0 python bench.py $X > gpurun_out/r02_c11_l0_cfg2.json 2>> gpurun_out/r02_c11.err
for f in gpurun_out/r02_c11_*cfg*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry']['tasks_per_evaluation'], r.get('reused_fraction'))"; done
tail -5 gpurun_out/r02_c11.err
