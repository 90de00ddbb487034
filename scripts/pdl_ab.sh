# GPU: the same library with and without programmatic dependent launch (B200_EMBED_PDL=0), interleaved
for r in 1 2 3; do for v in 0 1; do
B200_EMBED_PDL=$v python bench.py --workload ${W:-c2} --no-workloads --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/pdl$v.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/pdl$v.json") if l.startswith("{")][-1])
print("PDL=$v", round(d["value"]), "sent/s", round(d["ms_per_step"],3), "ms  e2e", round(d["e2e"]["value"]), d["clocks"]["sm_mhz"])
PY
done; done
