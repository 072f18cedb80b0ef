#!/bin/bash
# round 2, GPU call 1: tests, the driver's bench line, and the exploratory runs that size the light evaluations
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r02_pytest1.log
echo "pytest rc=$?" >> gpurun_out/r02_pytest1.log
python bench.py > gpurun_out/r02_bench_n1_a.json 2> gpurun_out/r02_bench_n1_a.err
X="--no-e2e --no-cpu --sustain-s 0"
python bench.py $X --iters 0 > gpurun_out/r02_x_iters0.json 2>> gpurun_out/r02_x.err
python bench.py $X --iters 1 > gpurun_out/r02_x_iters1.json 2>> gpurun_out/r02_x.err
python bench.py $X --config geman_lm > gpurun_out/r02_x_cfg3.json 2>> gpurun_out/r02_x.err
python bench.py $X --config geman_lm --iters 0 > gpurun_out/r02_x_cfg3_iters0.json 2>> gpurun_out/r02_x.err
python bench.py $X --config pyramid > gpurun_out/r02_x_cfg4.json 2>> gpurun_out/r02_x.err
python bench.py $X --grad-scale 0.125 > gpurun_out/r02_x_gs8.json 2>> gpurun_out/r02_x.err
python bench.py $X --grad-scale 0.015625 > gpurun_out/r02_x_gs64.json 2>> gpurun_out/r02_x.err
python bench.py $X --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_x_gs64_rot.json 2>> gpurun_out/r02_x.err
python bench.py $X --config geman_lm --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_x_cfg3_gs64_rot.json 2>> gpurun_out/r02_x.err
python bench.py --config geman_lm --total-queries 1024 > gpurun_out/r02_job_cfg3_1024.json 2>> gpurun_out/r02_x.err
tail -5 gpurun_out/r02_pytest1.log
