#!/bin/bash
# round 2, GPU call 16 (8 GPUs): the bench line at N = 8 (peer-store result gather, per-rank attribution), the e2e leg with
# host buffers interleaved over the NUMA nodes, cfg5 (16384-query job, strong scaling) at N = 8
mkdir -p gpurun_out
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
$T bench.py --gpus $N --no-cpu > gpurun_out/r02_c16_bench_n$N.json 2> gpurun_out/r02_c16_bench_n$N.err
$T bench.py --gpus $N --no-cpu --sustain-s 0 --no-every-row --no-adapter-e2e --host-numa interleave > gpurun_out/r02_c16_bench_n${N}_interleave.json 2>> gpurun_out/r02_c16_bench_n$N.err
$T bench.py --gpus $N --config cmu_slice --total-queries 16384 > gpurun_out/r02_c16_cfg5_n$N.json 2>> gpurun_out/r02_c16_bench_n$N.err
nvidia-smi topo -m > gpurun_out/r02_c16_topo_n$N.txt 2>&1; numactl -H >> gpurun_out/r02_c16_topo_n$N.txt 2>&1; lscpu | head -20 >> gpurun_out/r02_c16_topo_n$N.txt
for f in gpurun_out/r02_c16_*_n$N*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); pr=d.get('per_rank',{}); print(round(d['value']), round(d['ms_per_step'],3), pr.get('kernel_ms_min_median_max'), pr.get('step_over_slowest_kernel'), d.get('e2e',{}).get('value'), d.get('e2e',{}).get('h2d_gbs'), d.get('job',{}).get('wall_s_with_input_generation'))"; done
tail -n 3 gpurun_out/r02_c16_bench_n$N.err
