#!/bin/bash
# Runs on the GPU box (under gpurun): GPU parity tests by group, then a short bench. Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/box.txt 2>&1; nproc >> gpurun_out/box.txt; (cargo --version || echo "no cargo") >> gpurun_out/box.txt 2>&1
run() { name=$1; shift; echo "== $name"; timeout 1200 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -n ${TAILN:-12} gpurun_out/$name.log; }
run t_kernels python -m pytest tests/test_gpu_kernels.py -q --tb=short -m gpu -x
run t_e2e python -m pytest tests/test_gpu_e2e.py -q --tb=short -m gpu
run t_rest python -m pytest tests/test_gpu_search.py tests/test_zz_no_alloc.py -q --tb=short -m gpu
run smoke python __graft_entry__.py smoke
