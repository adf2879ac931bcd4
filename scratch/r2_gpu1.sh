#!/bin/bash
# round-2 GPU session 1: parity tests, quick decode timing, per-phase cycle breakdown, attention variant, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests_1.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2_gpu_tests_1.log
timeout 600 python scratch/quick_bench.py > gpurun_out/r2_quick_1.log 2>&1; echo "quick rc=$?"; tail -4 gpurun_out/r2_quick_1.log
B200LLM_ATT_PARTS=1 timeout 300 python scratch/quick_bench.py > gpurun_out/r2_quick_1_att1.log 2>&1; echo "quick att1 rc=$?"; head -1 gpurun_out/r2_quick_1_att1.log
B200LLM_ATT_PARTS=2 timeout 300 python scratch/quick_bench.py > gpurun_out/r2_quick_1_att2.log 2>&1; echo "quick att2 rc=$?"; head -1 gpurun_out/r2_quick_1_att2.log
timeout 300 python scratch/mega_dbg.py > gpurun_out/r2_mega_dbg_1.log 2>&1; echo "dbg rc=$?"; tail -25 gpurun_out/r2_mega_dbg_1.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_1.json 2> gpurun_out/r2_bench_1.err; echo "bench rc=$?"; cat gpurun_out/r2_bench_1.json
