#!/bin/bash
run() { timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/$1: /"; }
run "probe=0 poll=32"
B200LLM_LL_POLL_NS=100 run "probe=0 poll=100"
B200LLM_LL_POLL_NS=400 run "probe=0 poll=400"
B200LLM_LL_PROBE=3 run "probe=3 poll=32"
B200LLM_LL_PROBE=3 B200LLM_LL_POLL_NS=100 run "probe=3 poll=100"
B200LLM_LL_PROBE=3 B200LLM_LL_POLL_NS=400 run "probe=3 poll=400"
