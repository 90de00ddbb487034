#!/bin/bash
# Runs on the GPU box (under gpurun): the whole GPU suite by file, the smoke entry, then the default bench. Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/box.txt 2>&1; nproc >> gpurun_out/box.txt; (cargo --version || echo "no cargo") >> gpurun_out/box.txt 2>&1
run() { name=$1; shift; echo "== $name"; timeout 1200 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -n ${TAILN:-6} gpurun_out/$name.log; }
run t_all python -m pytest tests -q --tb=short -m gpu -x
run smoke python __graft_entry__.py smoke
if [ "${BENCH:-1}" = "1" ]; then run bench python bench.py; fi
