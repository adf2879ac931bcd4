#!/bin/bash
# GPU session: quick decode timing first (cheap, decides the kernel direction), then the whole GPU test suite, the bench line, an ncu capture
mkdir -p gpurun_out
for cfg in "32 1" "0 0" "200 0"; do
  set -- $cfg
  B200LLM_LL_POLL_NS=$1 B200LLM_LL_PROBE=$2 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/poll_ns=$1 probe=$2: /"
done
for ah in 0 16384 65536 131072; do
  B200LLM_MEGA_L2_AHEAD=$ah timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/l2_ahead=$ah: /"
done
for v in s3t6912 s2t11520 s4t4608; do
  B200LLM_LIB=$PWD/localai_b200/lib/libb200llm_$v.so timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/variant $v: /"
done
B200LLM_ATT_PARTS=1 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/att_parts=1: /"
B200LLM_ATT_PARTS=2 timeout 300 python scratch/quick_bench.py 2>&1 | grep "ms/step" | sed "s/^/att_parts=2: /"
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_gpu_tests_4.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|error|sigma|agreement" gpurun_out/r2_gpu_tests_4.log | tail -20
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_4.json 2> gpurun_out/r2_bench_4.err; echo "bench rc=$?"; cat gpurun_out/r2_bench_4.json; tail -3 gpurun_out/r2_bench_4.err
timeout 600 python scratch/prof_mega.py > gpurun_out/r2_prof_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 4 -c 1 -f -o gpurun_out/r2_mega_v9 python scratch/prof_mega.py > gpurun_out/r2_ncu_mega.log 2>&1
echo "ncu rc=$?"
