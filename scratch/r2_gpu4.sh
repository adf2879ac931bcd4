#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_engine_gpu.py -m gpu -x -q > gpurun_out/r2_gpu_tests_4.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r2_gpu_tests_4.log
for cfg in "32 1" "0 0" "200 0"; do
  set -- $cfg
  B200LLM_LL_POLL_NS=$1 B200LLM_LL_PROBE=$2 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/poll_ns=$1 probe=$2: /"
done
B200LLM_ATT_PARTS=1 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/att_parts=1: /"
B200LLM_ATT_PARTS=2 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/att_parts=2: /"
