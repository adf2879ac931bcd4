cd /root/repo
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "qgemm" 2>&1 | tail -15
