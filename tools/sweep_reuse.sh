#!/bin/bash
# Record reuse on / off on the bench workloads (value, ms per step, rows reused).  Usage: tools/sweep_reuse.sh <tag>
tag=${1:-reuse}
o=gpurun_out
small="--steps 5 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e"
for cfg in "--config gn_inlier" "--config geman_lm" "--config pyramid --points 8192" "--config geman_lm --queries 1" "--config geman_lm --queries 8"; do
    echo "== $cfg"
    timeout 300 python bench.py $small $cfg 2> $o/${tag}_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; e=r.get('every_row_gathered',{})
print('reuse: %.0f q/s  %.2f ms/step  kernel %.2f ms  contract frac %.2f  gathered %.0f GB/s (%.2f of peak)  rows reused %.3f' % (d['value'], d['ms_per_step'], r['kernel_ms'], r['frac'], r['gathered']['gbs'], r['gathered']['frac_of_peak'], r['gathered']['reused_fraction_of_rows']))
print('every row: %.0f q/s  %.2f ms/step  kernel %.2f ms  frac %.2f  identical %s' % (e.get('value',0), e.get('ms_per_step',0), e.get('kernel_ms',0), e.get('frac',0), e.get('results_bit_identical_to_default')))
"
done
