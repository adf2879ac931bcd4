#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
run() { timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/$1: /"; }
run "taper on"
B200LLM_MEGA_TAPER=0 run "taper off"
timeout 900 python -m pytest tests/test_tp_gpu.py tests/test_engine_gpu.py -m gpu -x -q -s -k "tensor_parallel or replica_pool" > gpurun_out/r2_gpu_tests_7.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/r2_gpu_tests_7.log
