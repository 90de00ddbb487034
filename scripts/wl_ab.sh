# GPU: one workload with two library builds, interleaved.  usage: W=c3 bash scripts/wl_ab.sh lib_a.so lib_b.so
for r in 1 2 3; do for l in "$@"; do
B200_EMBED_LIB=$PWD/$l python bench.py --workload ${W:-c3} --no-workloads --no-cpu-baseline --steps ${STEPS:-10} --warmup 3 > gpurun_out/wl.json 2>/dev/null
python - <<PY
import json
d=json.loads([x for x in open("gpurun_out/wl.json") if x.startswith("{")][-1])
k=d["kernels"]
print("${W:-c3} $l", round(d["value"]), "sent/s", round(d["ms_per_step"],3), "ms", d["clocks"]["sm_mhz"], {n: k[n]["ms_per_launch"] for n in ("attention","qkv_attention") if n in k})
PY
done; done
