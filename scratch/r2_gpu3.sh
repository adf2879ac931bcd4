#!/bin/bash
mkdir -p gpurun_out
for cfg in "32 1" "0 1" "0 0" "200 1" "200 0" "1000 1" "32 0"; do
  set -- $cfg
  B200LLM_LL_POLL_NS=$1 B200LLM_LL_PROBE=$2 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/poll_ns=$1 probe=$2: /"
done
