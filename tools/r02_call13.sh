#!/bin/bash
# round 2, GPU call 13 (2 GPUs): the default bench line at N = 2 (peer-store result gather), cfg5 job of 16384 queries at N = 2,
# cfg3 / cfg4 jobs of 1024 queries
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$T bench.py --gpus 2 > gpurun_out/r02_c13_bench_n2.json 2> gpurun_out/r02_c13_bench_n2.err
$T bench.py --gpus 2 --config cmu_slice --total-queries 16384 > gpurun_out/r02_c13_cfg5_n2.json 2> gpurun_out/r02_c13_cfg5_n2.err
$T bench.py --gpus 2 --config geman_lm --total-queries 1024 > gpurun_out/r02_c13_cfg3_1024_n2.json 2>> gpurun_out/r02_c13_cfg5_n2.err
$T bench.py --gpus 2 --config pyramid --total-queries 1024 > gpurun_out/r02_c13_cfg4_1024_n2.json 2>> gpurun_out/r02_c13_cfg5_n2.err
for f in gpurun_out/r02_c13_*.json; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); print(round(d['value']), d['ms_per_step'], d.get('per_rank',{}).get('kernel_ms'), d.get('e2e',{}).get('value'), d.get('job',{}).get('wall_s_with_input_generation'))"; done
tail -3 gpurun_out/r02_c13_*.err
