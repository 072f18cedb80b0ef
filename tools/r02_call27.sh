#!/bin/bash
# round 2, GPU call 27: record layout [c, tag | Sxx, Sxy | Syy, bx | by, pad] (N) against the committed [c, Sxx, Sxy, Syy | bx, by, tag, pad] (A)
mkdir -p gpurun_out
FMPNP_LIB=$PWD/gpurun_in/lib_N.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_random_shapes.py tests/test_gpu_edge_and_scale.py tests/test_gpu_fullsize_parity.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -n 3
X="--no-e2e --no-cpu --sustain-s 0"
for V in N A N A; do
  export FMPNP_LIB=$PWD/gpurun_in/lib_$V.so
  python bench.py $X >> gpurun_out/r02_c27_${V}_cfg2.json 2>> gpurun_out/r02_c27.err
  python bench.py $X --no-every-row --config geman_lm >> gpurun_out/r02_c27_${V}_cfg3.json 2>> gpurun_out/r02_c27.err
  python bench.py $X --no-every-row --config pyramid >> gpurun_out/r02_c27_${V}_cfg4.json 2>> gpurun_out/r02_c27.err
  python bench.py $X --no-every-row --iters 100 >> gpurun_out/r02_c27_${V}_it100.json 2>> gpurun_out/r02_c27.err
done
for f in gpurun_out/r02_c27_*.json; do echo $f; python -c "
import json,sys
for l in open('$f').read().strip().splitlines():
    d=json.loads(l); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r.get('every_row_kernel_ms'))"; done
