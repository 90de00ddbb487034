#!/bin/bash
# On the GPU box, ONE profiler pass per call (after the same command has exited 0 without the profiler):
#   gpu_profile.sh list                 launch list: device time of every kernel of a short bench run
#   gpu_profile.sh full [regex] [skip] [count]   ncu --set full of the kernels matching regex
mkdir -p gpurun_out
CMD="python bench.py --workload ${WL:-c2} --steps 2 --warmup 1 --no-cpu-baseline --no-workloads"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
if [ "$1" = "list" ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "launch list exit $?"
else
  ncu --set full --clock-control none --import-source on -k regex:"${2:-gemm_tc_kernel|ffn_fused|attention_tc|embed_layernorm|mean_pool}" -s ${3:-30} -c ${4:-8} -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  echo "full capture exit $?"; tail -3 gpurun_out/ncu_full.log
fi
