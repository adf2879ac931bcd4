#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 600 python bench.py --steps 8 --warmup 3 --no-extras > gpurun_out/r2_bench_small_v14.json 2>/dev/null &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_mega_kernel|dequant_rows_kernel|argmax_kernel" -c 200 --csv --log-file gpurun_out/r2_bench_launches_v14.csv python bench.py --steps 8 --warmup 3 --no-extras > gpurun_out/r2_ncu_bench_v14.log 2>&1
echo "ncu launches rc=$?"
