#!/bin/bash
# round 2, GPU call 23: code-size variants of the task kernel: A = committed, B = one call site for the per-point phase of a task,
# C = B + stream_list as a function of its own
mkdir -p gpurun_out
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for V in A B C A B C; do
  if [ $V = A ]; then unset FMPNP_LIB; else export FMPNP_LIB=$PWD/gpurun_in/lib_$V.so; fi
  python bench.py $X >> gpurun_out/r02_c23_${V}_cfg2.json 2>> gpurun_out/r02_c23.err
  python bench.py $X --config geman_lm >> gpurun_out/r02_c23_${V}_cfg3.json 2>> gpurun_out/r02_c23.err
  python bench.py $X --config pyramid >> gpurun_out/r02_c23_${V}_cfg4.json 2>> gpurun_out/r02_c23.err
done
for V in B C; do
  FMPNP_LIB=$PWD/gpurun_in/lib_$V.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_edge_and_scale.py -m gpu -x -q 2>&1 | tail -n 2 > gpurun_out/r02_c23_pytest_$V.log
  cat gpurun_out/r02_c23_pytest_$V.log
done
for f in gpurun_out/r02_c23_*.json; do echo $f; python -c "
import json,sys
for l in open('$f').read().strip().splitlines():
    d=json.loads(l); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3))"; done
