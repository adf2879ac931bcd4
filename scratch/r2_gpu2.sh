#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests_2.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2_gpu_tests_2.log
timeout 600 python scratch/quick_bench.py > gpurun_out/r2_quick_2.log 2>&1; echo "quick rc=$?"; tail -4 gpurun_out/r2_quick_2.log
B200LLM_ATT_PARTS=1 timeout 300 python scratch/quick_bench.py > gpurun_out/r2_quick_2_att1.log 2>&1; echo "quick att1 rc=$?"; head -1 gpurun_out/r2_quick_2_att1.log
B200LLM_NO_DIST_QUANT=1 timeout 300 python scratch/quick_bench.py > gpurun_out/r2_quick_2_nodq.log 2>&1; echo "quick nodq rc=$?"; head -1 gpurun_out/r2_quick_2_nodq.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_2.json 2> gpurun_out/r2_bench_2.err; echo "bench rc=$?"; cat gpurun_out/r2_bench_2.json
