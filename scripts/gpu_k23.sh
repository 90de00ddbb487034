#!/bin/bash
# GPU box: K23 parity tests, the fused-schedule end-to-end tests, then a short bench (headline config only).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_e2e.py -q --tb=short -m gpu -x -k "k23 or fused or config5 or scale" 2>&1 | tail -8
timeout 600 python -m pytest tests/test_gpu_configs.py -q --tb=short -m gpu -x -k "config5 or config2" 2>&1 | tail -4
for ops in ${OPS:-bf16}; do
timeout 300 python bench.py --steps ${STEPS:-10} --warmup 3 --no-workloads --no-cpu-baseline --operands $ops > gpurun_out/k23_bench_$ops.json 2> gpurun_out/k23_bench_$ops.err || tail -5 gpurun_out/k23_bench_$ops.err
python - <<PY
import json
d=json.load(open("gpurun_out/k23_bench_$ops.json"))
print("$ops", round(d["value"]), "sent/s", round(d["ms_per_step"],3), "ms/step", d["clocks"]["sm_mhz"], d["clocks"]["reasons"], "model frac", d["model_frac_of_sustained_bf16_peak"])
for k,v in d["kernels"].items(): print("  ", k, v["ms_per_launch"], v["share_of_step"], v["frac"])
PY
done
