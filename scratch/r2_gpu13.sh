#!/bin/bash
mkdir -p gpurun_out
run() { timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/$1: /"; }
run "default"
B200LLM_LL_PROBE=2 run "hint"
timeout 300 python scratch/mega_dbg.py > gpurun_out/r2_mega_dbg_13.log 2>&1; tail -26 gpurun_out/r2_mega_dbg_13.log
