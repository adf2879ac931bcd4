cd /root/repo
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "qgemm" 2>&1 | tail -3
timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
B200LLM_NO_GRAPH=1 STEPS=2 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/b64q_launches.csv python scratch/b64_bench.py > gpurun_out/b64q_ncu.log 2>&1; echo rc=$?
