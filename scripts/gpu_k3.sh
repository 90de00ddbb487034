#!/bin/bash
# GPU box: K3 (attention_tc) parity tests and the workloads that use it (c3 ragged 8..512, c4 BERT-base), A/B against a baseline build.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_e2e.py tests/test_gpu_configs.py -q --tb=short -m gpu -x -k "k3_attention or config3 or config4 or xlmr or config1 or golden" 2>&1 | tail -6
for lib in ${LIBS:-codename-symbiont_b200/libb200_embed_base.so codename-symbiont_b200/libb200_embed.so}; do
  for w in c3 c4; do
    B200_EMBED_LIB=$PWD/$lib python bench.py --workload $w --steps 3 --warmup 3 --no-workloads --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$lib $w', round(d['value']), 'sent/s', round(d['ms_per_step'],2), 'ms', d['clocks']['sm_mhz'], 'attention', d['kernels']['attention']['ms_per_launch'], d['kernels']['attention']['share_of_step'])"
  done
done
