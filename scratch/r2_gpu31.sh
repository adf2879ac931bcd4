cd /root/repo
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "qgemm" 2>&1 | tail -3
timeout 300 python scratch/qgemm_dbg.py 2>&1 | grep "type\|CTA "
timeout 300 python scratch/qg_kbench.py 2>&1 | tail -6
