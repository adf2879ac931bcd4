cd /root/repo
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "qgemm and 28672" 2>&1 | tail -2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qgemm_kernel -c 2 -f -o gpurun_out/qgemm_v1 python -m pytest tests/test_parity_gpu.py -x -q -k "qgemm and 28672" > gpurun_out/qgemm_ncu.log 2>&1; echo rc=$?
