#!/bin/bash
for st in 0 1 2 3; do for ld in 0 1; do
  B200LLM_TP_ST=$st B200LLM_TP_LD=$ld timeout 300 python scratch/tp_bench.py llama3-8b 2 2>&1 | grep "fused" | sed "s/^/st=$st ld=$ld: /"
done; done
