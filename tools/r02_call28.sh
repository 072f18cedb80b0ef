#!/bin/bash
# round 2, GPU call 28 (N GPUs): BASELINE configs 3 and 4 at their stated 1024 queries, sharded over N ranks (job mode, strong scaling)
mkdir -p gpurun_out
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
$T bench.py --gpus $N --config geman_lm --total-queries 1024 > gpurun_out/r02_c28_cfg3_1024_n$N.json 2> gpurun_out/r02_c28_n$N.err
$T bench.py --gpus $N --config pyramid --total-queries 1024 > gpurun_out/r02_c28_cfg4_1024_n$N.json 2>> gpurun_out/r02_c28_n$N.err
for f in gpurun_out/r02_c28_*_n$N.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],3), d['job']['chunks_per_rank'], d['job']['queries_finished'])"; done
