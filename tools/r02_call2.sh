#!/bin/bash
# round 2, GPU call 2: the no-allocation entry point (alone in a fresh process, then after every other test), smoke, bench line
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r02_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest2.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_pytest2.log
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_smoke2.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r02_smoke2.log
python bench.py > gpurun_out/r02_bench_n1_b.json 2> gpurun_out/r02_bench_n1_b.err
tail -3 gpurun_out/r02_noalloc_alone.log gpurun_out/r02_pytest2.log gpurun_out/r02_smoke2.log
head -c 600 gpurun_out/r02_bench_n1_b.json
