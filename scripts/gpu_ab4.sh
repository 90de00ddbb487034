#!/bin/bash
# A/B of environment switches on the default bench (c2); prints per-kind ms per launch.
run() {
  echo "== $*"
  env "$@" python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | grep "^{" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms',round(d['ms_per_step'],2), d['clocks']['sm_mhz'], {k:v['ms_per_launch'] for k,v in d['kernels'].items()})"
}
run X=0
run B200_EMBED_QKV_ATTN=1
