#!/bin/bash
# round 2, GPU call 22: the big phases as functions of their own (__noinline__): parity subset, timings
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_edge_and_scale.py tests/test_gpu_fullsize_parity.py -m gpu -x -q > gpurun_out/r02_c22_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c22_pytest.log
tail -n 3 gpurun_out/r02_c22_pytest.log
X="--no-e2e --no-cpu --sustain-s 0"
for V in new old; do
  if [ $V = old ]; then export FMPNP_LIB=$PWD/gpurun_in/libfmpnp_inlined.so; else unset FMPNP_LIB; fi
  python bench.py $X > gpurun_out/r02_c22_${V}_cfg2.json 2>> gpurun_out/r02_c22.err
  python bench.py $X --no-every-row --config geman_lm > gpurun_out/r02_c22_${V}_cfg3.json 2>> gpurun_out/r02_c22.err
  python bench.py $X --config pyramid > gpurun_out/r02_c22_${V}_cfg4.json 2>> gpurun_out/r02_c22.err
  python bench.py $X --no-every-row --queries 1 --config geman_lm > gpurun_out/r02_c22_${V}_q1.json 2>> gpurun_out/r02_c22.err
  python bench.py $X --no-every-row --queries 32 --config geman_lm > gpurun_out/r02_c22_${V}_q32.json 2>> gpurun_out/r02_c22.err
  python bench.py $X --no-every-row --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c22_${V}_gs64rot.json 2>> gpurun_out/r02_c22.err
done
for f in gpurun_out/r02_c22_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r.get('every_row_kernel_ms'))"; done
tail -n 3 gpurun_out/r02_c22.err
