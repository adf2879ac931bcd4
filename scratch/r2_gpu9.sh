set -x
cd /root/repo
timeout 300 python scratch/b64_bench.py > gpurun_out/b64_plain.log 2>&1; echo rc=$?
B200LLM_NO_GRAPH=1 STEPS=2 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/b64_launches.csv python scratch/b64_bench.py > gpurun_out/b64_ncu.log 2>&1; echo rc=$?
tail -3 gpurun_out/b64_plain.log
