#!/bin/bash
# round 2, GPU call 25: ncu --set full of a launch dominated by light evaluations (100 iterations), light route on
mkdir -p gpurun_out
# default policy (the light route is on for this workload)
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --sustain-s 0 --no-every-row --iters 100"
$CMD > gpurun_out/r02_c25_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:refine_tasks -s 2 -c 1 -o gpurun_out/r02_c25_light100 $CMD > gpurun_out/r02_c25_ncu.log 2>&1
echo "rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/r02_c25_plain.log').read().strip().splitlines()[-1]); print(d['value'], d['roofline']['kernel_ms'])"
