#!/bin/bash
# round 2, GPU call 12: experiment -- larger fixed reduction ranges (64 / 128 points instead of 32): parity subset and timings
mkdir -p gpurun_out
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for RP in 128 64 32; do
  if [ $RP = 32 ]; then unset FMPNP_LIB; else export FMPNP_LIB=$PWD/gpurun_in/libfmpnp_rp$RP.so; fi
  python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py -m gpu -x -q 2>&1 | tail -2 > gpurun_out/r02_c12_rp${RP}_pytest.log
  python bench.py $X > gpurun_out/r02_c12_rp${RP}_cfg2.json 2>> gpurun_out/r02_c12.err
  python bench.py $X --config geman_lm > gpurun_out/r02_c12_rp${RP}_cfg3.json 2>> gpurun_out/r02_c12.err
  python bench.py $X --config pyramid > gpurun_out/r02_c12_rp${RP}_cfg4.json 2>> gpurun_out/r02_c12.err
  python bench.py $X --queries 1 --config geman_lm > gpurun_out/r02_c12_rp${RP}_q1.json 2>> gpurun_out/r02_c12.err
  python bench.py $X --queries 32 --config geman_lm > gpurun_out/r02_c12_rp${RP}_q32.json 2>> gpurun_out/r02_c12.err
done
tail -2 gpurun_out/r02_c12_rp*_pytest.log
for f in gpurun_out/r02_c12_rp*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry']['tasks_per_evaluation'])"; done
tail -5 gpurun_out/r02_c12.err
