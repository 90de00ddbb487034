#!/bin/bash
for e in "X=0" "B200_EMBED_LN_LEAN_OFF=1"; do
  echo "== $e"
  env $e python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "^{" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms',round(d['ms_per_step'],2), d.get('model_frac_of_sustained_bf16_peak'), {k:v['ms_per_launch'] for k,v in d['kernels'].items()})"
done
