#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests_v15.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2_gpu_tests_v15.log
timeout 1200 python bench.py > gpurun_out/r2_bench_v15.json 2> gpurun_out/r2_bench_v15.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_v15.json')); print({k:d[k] for k in ('value','ms_per_step','parity_checked','steps','warmup')}); print(d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'] if d['cpu_baseline'] else None); print(json.dumps(d['extra'])[:3000])"
