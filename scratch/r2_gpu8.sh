#!/bin/bash
mkdir -p gpurun_out
free -g | head -2; nproc; df -h /dev/shm | tail -1
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench rc=$?"
cat gpurun_out/r2_bench_n2.json; tail -15 gpurun_out/r2_bench_n2.err
