#!/bin/bash
# GPU: BERT-base workload (c4) with the LayerNorm GEMM's cluster shapes: 4 x 192 columns (shipped) vs 3 x 256 (B200_EMBED_LN_CL3=1)
for r in 1 2; do for e in "X=0" "B200_EMBED_LN_CL3=1"; do
  echo "== $e"
  env $e timeout 300 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline --no-workloads 2>&1 | grep "^{" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms',round(d['ms_per_step'],2), d.get('model_frac_of_sustained_bf16_peak'), d['clocks']['sm_mhz'], {k:v['ms_per_launch'] for k,v in d['kernels'].items()})"
done; done
