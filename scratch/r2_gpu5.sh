#!/bin/bash
mkdir -p gpurun_out
run() { timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/$1: /"; }
run "default (probe=1 poll=32 l2=0)"
B200LLM_LL_PROBE=0 run "probe=0"
B200LLM_LL_POLL_NS=200 run "poll=200"
B200LLM_LL_POLL_NS=0 run "poll=0"
B200LLM_MEGA_L2_AHEAD=16384 run "l2_ahead=16K"
B200LLM_MEGA_L2_AHEAD=32768 run "l2_ahead=32K"
B200LLM_NO_DIST_QUANT=1 run "no dist quant"
B200LLM_ATT_PARTS=2 run "att_parts=2"
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_gpu_tests_5.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|error|sigma|agreement" gpurun_out/r2_gpu_tests_5.log | tail -12
timeout 300 python scratch/mega_dbg.py > gpurun_out/r2_mega_dbg_5.log 2>&1; tail -22 gpurun_out/r2_mega_dbg_5.log
