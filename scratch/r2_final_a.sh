#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_final.json')); print({k:d[k] for k in ('value','ms_per_step','parity_checked','steps','warmup')}); print(d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'] if d['cpu_baseline'] else None); print({k:(v.get('tok_s') or v.get('per_call_16',{}).get('seq_s')) for k,v in d['extra'].items()})"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null; echo "ref rc=$?"; cut -c1-400 gpurun_out/r2_bench_reference.json
timeout 600 python bench.py --steps 8 --warmup 3 --no-extras > gpurun_out/r2_bench_small.json 2>/dev/null &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_mega_kernel|dequant_rows_kernel|argmax_kernel" -c 200 --csv --log-file gpurun_out/r2_bench_launches_final.csv python bench.py --steps 8 --warmup 3 --no-extras > gpurun_out/r2_ncu_bench.log 2>&1
echo "ncu launches rc=$?"
