#!/bin/bash
# round 2, GPU call 18: does a light evaluation's time depend on the CTAs that share its SM?  One CTA per SM x 148 queries against
# two per SM x 296 queries (same work per CTA); fmpnp launch list; compute-sanitizer memcheck on the small all-paths case
mkdir -p gpurun_out
X="--no-e2e --no-cpu --sustain-s 0 --no-every-row"
for IT in 0 20; do
  FMPNP_LIGHT=1 FMPNP_CTAS_PER_SM=1 python bench.py $X --queries 148 --iters $IT > gpurun_out/r02_c18_ps1_q148_it$IT.json 2>> gpurun_out/r02_c18.err
  FMPNP_LIGHT=1 python bench.py $X --queries 296 --iters $IT > gpurun_out/r02_c18_ps2_q296_it$IT.json 2>> gpurun_out/r02_c18.err
  FMPNP_LIGHT=1 python bench.py $X --queries 148 --iters $IT > gpurun_out/r02_c18_ps2_q148_it$IT.json 2>> gpurun_out/r02_c18.err
done
for f in gpurun_out/r02_c18_ps*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d['roofline']; print(round(d['value']), round(r['kernel_ms'],3), r['geometry'])"; done
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --sustain-s 0"
$CMD > gpurun_out/r02_c18_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:tasks_kernel -c 60 --csv --log-file gpurun_out/r02_c18_launches_fmpnp.csv $CMD > gpurun_out/r02_c18_ncu_list.log 2>&1
echo "ncu rc=$?"
