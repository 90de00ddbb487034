#!/bin/bash
for env in "X=0" "B200_EMBED_GEMM_DBG=1" "B200_EMBED_GEMM_DBG=2" "B200_EMBED_GEMM_DBG=3"; do
  echo "== $env"
  env $env python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | grep "^{" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms',round(d['ms_per_step'],2), {k:v['ms_per_launch'] for k,v in d['kernels'].items() if 'gemm' in k or k=='attention'})"
done
