set -x
cd /root/repo
timeout -s KILL 600 python -m pytest tests/test_parity_gpu.py -x -q -k "fused_and_mega" 2>&1 | tail -3
timeout -s KILL 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout -s KILL 300 python scratch/quick_bench.py 2>&1 | grep "ms/step"
timeout -s KILL 300 python scratch/mega_dbg.py 2>&1 | tail -16
