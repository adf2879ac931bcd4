#!/bin/bash
# full GPU suite, the bench command (plain, then its ncu launch list), and an ncu --set full capture of the decode kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_gpu_tests_14.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|error" gpurun_out/r2_gpu_tests_14.log | tail -5
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_14.json 2> gpurun_out/r2_bench_14.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_14.json')); print({k:d[k] for k in ('value','ms_per_step','parity_checked')}); print(d['roofline']['frac'], d['e2e']['value']); print({k:(v.get('tok_s') or v.get('per_call_16',{}).get('seq_s')) for k,v in d['extra'].items()})"
timeout 900 python bench.py --steps 4 --warmup 3 --no-extras > gpurun_out/r2_bench_small.json 2>/dev/null &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --steps 4 --warmup 3 --no-extras > gpurun_out/r2_ncu_bench.log 2>&1
echo "ncu launches rc=$?"
timeout 600 python scratch/prof_mega.py > gpurun_out/r2_prof_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 4 -c 1 -f -o gpurun_out/r2_mega_v10 python scratch/prof_mega.py > gpurun_out/r2_ncu_mega.log 2>&1
echo "ncu full rc=$?"
