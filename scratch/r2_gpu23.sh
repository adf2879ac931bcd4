cd /root/repo
timeout 300 python scratch/qgemm_dbg.py 2>&1 | tail -20
