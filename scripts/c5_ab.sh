L=codename-symbiont_b200/libb200_embed
for r in 1 2 3; do for v in _nb ""; do
B200_EMBED_LIB=$PWD/${L}${v}.so python bench.py --workload c5 --no-workloads --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/c5$v.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/c5$v.json") if l.startswith("{")][-1])
print("variant '$v'", round(d["value"]), "sent/s", round(d["ms_per_step"],3), "ms; qkv", d["kernels"]["qkv_attention"]["ms_per_launch"])
PY
done; done
