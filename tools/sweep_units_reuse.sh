#!/bin/bash
# Tasks per evaluation (FMPNP_UNITS) with record reuse: the light evaluations pay per task.  Usage: tools/sweep_units_reuse.sh
small="--steps 10 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e"
for cfg in "--config gn_inlier" "--config geman_lm" "--config pyramid --points 8192"; do
  for u in 0 1 2 4 8 16; do
    echo -n "== $cfg FMPNP_UNITS=$u: "
    FMPNP_UNITS=$u timeout 300 python bench.py $small $cfg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; e=r.get('every_row_gathered',{})
print('%.0f q/s kernel %.3f ms | every row %.0f q/s %.2f ms | U=%s' % (d['value'], r['kernel_ms'], e.get('value',0), e.get('kernel_ms',0), r['geometry']['tasks_per_evaluation']))
"
  done
done
