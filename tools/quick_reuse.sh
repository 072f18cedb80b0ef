small="--steps 10 --warmup 3 --no-e2e --no-cpu --no-adapter-e2e"
for cfg in "--config gn_inlier" "--config geman_lm" "--config pyramid --points 8192" "--config geman_lm --queries 32" "--config geman_lm --queries 1"; do
    echo -n "== $cfg: "
    timeout 300 python bench.py $small $cfg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; e=r.get('every_row_gathered',{})
print('%.0f q/s kernel %.3f ms | every row %.0f q/s %.2f ms frac %.3f | U=%s same=%s' % (d['value'], r['kernel_ms'], e.get('value',0), e.get('kernel_ms',0), e.get('frac',0), r['geometry']['tasks_per_evaluation'], e.get('results_bit_identical_to_default')))
"
done
