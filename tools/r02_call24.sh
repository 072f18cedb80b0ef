#!/bin/bash
# round 2, GPU call 24: final validation of the committed build: suite (alone-first no-alloc test, then everything), smoke, the driver's
# two bench commands, job lines of configs 3 / 4 / 5 at N = 1
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02_c24_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02_c24_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c24_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_c24_pytest.log
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_c24_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r02_c24_smoke.log
tail -n 2 gpurun_out/r02_c24_noalloc_alone.log gpurun_out/r02_c24_pytest.log gpurun_out/r02_c24_smoke.log
python bench.py --impl reference > gpurun_out/r02_c24_bench_ref.json 2> gpurun_out/r02_c24_bench.err
python bench.py > gpurun_out/r02_c24_bench_n1.json 2>> gpurun_out/r02_c24_bench.err
python bench.py --config geman_lm --total-queries 1024 > gpurun_out/r02_c24_cfg3_1024_n1.json 2>> gpurun_out/r02_c24_bench.err
python bench.py --config pyramid --total-queries 1024 > gpurun_out/r02_c24_cfg4_1024_n1.json 2>> gpurun_out/r02_c24_bench.err
python bench.py --config cmu_slice --total-queries 16384 > gpurun_out/r02_c24_cfg5_n1.json 2>> gpurun_out/r02_c24_bench.err
for f in gpurun_out/r02_c24_bench_ref.json gpurun_out/r02_c24_bench_n1.json gpurun_out/r02_c24_cfg*_n1.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d.get('roofline',{}); print(round(d['value'],2), round(d['ms_per_step'],3), r.get('frac'), r.get('every_row_frac'), d.get('e2e',{}).get('value'), d.get('parity_vs_cpu_arm',{}).get('within_north_star'), d.get('job',{}).get('wall_s_with_input_generation'))"; done
tail -n 3 gpurun_out/r02_c24_bench.err
