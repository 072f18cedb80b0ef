#!/bin/bash
# round 2, GPU call 15: dynamic light / task decision (previous evaluation's misses): suite, moving workloads, launch list
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c15_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c15_pytest.log
tail -n 3 gpurun_out/r02_c15_pytest.log
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
python bench.py $X > gpurun_out/r02_c15_x_cfg2.json 2>> gpurun_out/r02_c15_x.err
python bench.py $X --grad-scale 0.125 > gpurun_out/r02_c15_x_gs8.json 2>> gpurun_out/r02_c15_x.err
python bench.py $X --grad-scale 0.015625 > gpurun_out/r02_c15_x_gs64.json 2>> gpurun_out/r02_c15_x.err
python bench.py $X --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c15_x_gs64_rot.json 2>> gpurun_out/r02_c15_x.err
FMPNP_LIGHT=0 python bench.py $X --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c15_x_gs64_rot_l0.json 2>> gpurun_out/r02_c15_x.err
FMPNP_LIGHT=0 python bench.py $X --grad-scale 0.125 > gpurun_out/r02_c15_x_gs8_l0.json 2>> gpurun_out/r02_c15_x.err
FMPNP_LIGHT_MISS_DEN=256 python bench.py $X > gpurun_out/r02_c15_x_cfg2_den256.json 2>> gpurun_out/r02_c15_x.err
FMPNP_LIGHT_MISS_DEN=16 python bench.py $X --grad-scale 0.125 > gpurun_out/r02_c15_x_gs8_den16.json 2>> gpurun_out/r02_c15_x.err
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --sustain-s 0"
$CMD > gpurun_out/r02_c15_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fmpnp -c 60 --csv --log-file gpurun_out/r02_c15_launches_fmpnp.csv $CMD > gpurun_out/r02_c15_ncu_list.log 2>&1
for f in gpurun_out/r02_c15_x_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r.get('reused_fraction'), r.get('frac'))"; done
