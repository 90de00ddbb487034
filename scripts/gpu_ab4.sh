#!/bin/bash
for env in "X=0" "B200_EMBED_GEMM_WIDE_STORE=1" "B200_EMBED_GEMM_WIDE_STORE=1 B200_EMBED_QKV_BN128=1"; do
  echo "== $env"
  env $env python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | grep "^{" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms',round(d['ms_per_step'],2), d['clocks']['sm_mhz'], {k:v['ms_per_launch'] for k,v in d['kernels'].items() if 'gemm' in k or k in ('attention','ffn_fused')})"
done
env B200_EMBED_GEMM_WIDE_STORE=1 python -m pytest tests/test_gpu_e2e.py -q -m gpu -k "config1 or golden_minilm" 2>&1 | tail -1
