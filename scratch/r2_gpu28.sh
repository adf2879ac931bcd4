cd /root/repo
timeout 300 python scratch/qg_kbench.py 2>&1 | tail -7
