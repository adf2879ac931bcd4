cd /root/repo
timeout 300 python scratch/qg_kbench.py 2>&1 | tail -6
NB=8 timeout 300 python scratch/qg_kbench.py 2>&1 | tail -6
timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
NB=8 timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
NB=8 B200LLM_NO_QGEMM=1 timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
NB=2 timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
NB=2 B200LLM_NO_QGEMM=1 timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
