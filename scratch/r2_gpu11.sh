#!/bin/bash
B200LLM_MEGA_DBG=1 timeout 900 python scratch/tp_bench.py llama3-70b 2 2>&1 | grep -v Warning | tail -10
