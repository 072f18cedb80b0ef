#!/bin/bash
# round 2, GPU call 5: ncu --set full of one refinement launch with the light route on (cfg2)
mkdir -p gpurun_out
export FMPNP_LIGHT=1
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --sustain-s 0 --no-every-row"
$CMD > gpurun_out/r02_c5_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:refine_tasks -s 2 -c 1 -o gpurun_out/r02_c5_light $CMD > gpurun_out/r02_c5_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/r02_c5_ncu.log
