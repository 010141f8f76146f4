for be in 16777216 33554432 67108864; do
  echo "== BLOCK_EDGES $be"
  GNNCCA_BLOCK_EDGES=$be python bench.py --workload cfg4 --no-cpu-baseline --no-attach --steps 3 --warmup 3 --window-seconds 1.0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1])
print('ms_per_step', d['ms_per_step'], 'value', d['value'])
k=d['roofline'].get('kernels',{})
for n,v in k.items(): print('  ',n, v.get('ms'), v.get('frac'), v.get('contract_frac'))
"
done
