#!/bin/bash
# On the GPU box: launch list (device time of every kernel) and one full ncu capture of the GEMM kernels.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${1:-gemm_tc_kernel}" -s ${2:-40} -c ${3:-4} -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; tail -3 gpurun_out/ncu_full.log
