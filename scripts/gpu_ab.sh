#!/bin/bash
# A/B of GEMM schedules on the c2 bench (short runs)
for env in "B200_EMBED_GEMM_RESIDENT_A=1" "B200_EMBED_GEMM_RESIDENT_A=0" "B200_EMBED_GEMM_RESIDENT_A=0 B200_EMBED_GEMM_PAIR=1"; do
  echo "== $env"
  env $env python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms',round(d['ms_per_step'],2), {k:v['ms_per_launch'] for k,v in d['kernels'].items() if 'gemm' in k or k=='attention'})"
done
