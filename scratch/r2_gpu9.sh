#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scratch/tp_bench.py llama3-8b 2 2>&1 | grep -v Warning | tail -8
B200LLM_MEGA_DBG=1 timeout 600 python scratch/tp_bench.py llama3-8b 2 2>&1 | grep -v Warning | tail -10
