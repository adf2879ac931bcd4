#!/bin/bash
mkdir -p gpurun_out
run() { timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/$1: /"; }
run "default (probe=2 hint)"
B200LLM_LL_PROBE=0 run "probe=0"
B200LLM_LL_PROBE=2 B200LLM_LL_POLL_NS=100 run "hint poll=100"
B200LLM_LL_PROBE=2 B200LLM_ATT_PARTS=2 run "hint att_parts=2"
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_engine_gpu.py -m gpu -x -q > gpurun_out/r2_gpu_tests_6.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2_gpu_tests_6.log
timeout 300 python scratch/mega_dbg.py > gpurun_out/r2_mega_dbg_6.log 2>&1; tail -16 gpurun_out/r2_mega_dbg_6.log
B200LLM_LL_PROBE=0 timeout 300 python scratch/mega_dbg.py > gpurun_out/r2_mega_dbg_6_p0.log 2>&1; tail -16 gpurun_out/r2_mega_dbg_6_p0.log
