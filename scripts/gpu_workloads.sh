#!/bin/bash
# Short bench of every workload (c1..c4) -> gpurun_out/workloads.log; one line of key numbers each.
mkdir -p gpurun_out; : > gpurun_out/workloads.log
for w in c1 c3 c4 c5 c2; do
  python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | grep "^{" | tail -1 >> gpurun_out/workloads.log
done
python - <<'PY'
import json
for l in open("gpurun_out/workloads.log"):
    d = json.loads(l)
    print(d["config"]["workload"][:60], "| value", round(d["value"]), "| e2e", round(d["e2e"]["value"]), "| ms", round(d["ms_per_step"], 2),
          "| model", d.get("model_frac_of_sustained_bf16_peak"), "|", {k: v["ms_per_launch"] for k, v in d["kernels"].items()})
PY
