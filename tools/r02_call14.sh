#!/bin/bash
# round 2, GPU call 14: full suite + smoke on the 128-point-range build, gather ceilings incl. the TMA bulk variant, the
# driver's bench line, moving workloads, launch list + full ncu capture of the default command
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c14_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c14_pytest.log
tail -n 3 gpurun_out/r02_c14_pytest.log
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_c14_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r02_c14_smoke.log
tail -n 2 gpurun_out/r02_c14_smoke.log
for RB in 4096 2048 512; do timeout 200 tools/gather_peak $RB 32 1 >> gpurun_out/r02_c14_gather_peak.txt 2>&1; done
timeout 300 tools/gather_peak 4096 32 16 >> gpurun_out/r02_c14_gather_peak.txt 2>&1
cat gpurun_out/r02_c14_gather_peak.txt
python bench.py > gpurun_out/r02_c14_bench_n1.json 2> gpurun_out/r02_c14_bench_n1.err
python bench.py --impl reference > gpurun_out/r02_c14_bench_ref.json 2>> gpurun_out/r02_c14_bench_n1.err
X="--no-e2e --no-cpu --sustain-s 0"
python bench.py $X --grad-scale 0.125 > gpurun_out/r02_c14_x_gs8.json 2>> gpurun_out/r02_c14_x.err
python bench.py $X --grad-scale 0.015625 > gpurun_out/r02_c14_x_gs64.json 2>> gpurun_out/r02_c14_x.err
python bench.py $X --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c14_x_gs64_rot.json 2>> gpurun_out/r02_c14_x.err
python bench.py $X --config geman_lm > gpurun_out/r02_c14_x_cfg3.json 2>> gpurun_out/r02_c14_x.err
python bench.py $X --config geman_lm --grad-scale 0.015625 --rot-deg 1 3 > gpurun_out/r02_c14_x_cfg3_gs64_rot.json 2>> gpurun_out/r02_c14_x.err
python bench.py $X --config pyramid > gpurun_out/r02_c14_x_cfg4.json 2>> gpurun_out/r02_c14_x.err
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --sustain-s 0"
$CMD > gpurun_out/r02_c14_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_c14_launches.csv $CMD > gpurun_out/r02_c14_ncu_list.log 2>&1
$CMD > gpurun_out/r02_c14_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:refine_tasks -s 2 -c 1 -o gpurun_out/r02_c14_cfg2_full $CMD > gpurun_out/r02_c14_ncu_full.log 2>&1
echo "ncu rc=$?"
for f in gpurun_out/r02_c14_bench_n1.json gpurun_out/r02_c14_x_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r.get('reused_fraction'), r.get('frac'), r.get('every_row_frac'), d.get('e2e',{}).get('value'))"; done
