#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 300 python scratch/qg_kbench.py > gpurun_out/r2_qg_kbench.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qgemm_kernel -s 40 -c 4 -f -o gpurun_out/r2_qgemm_v2 python scratch/qg_kbench.py > gpurun_out/r2_ncu_qgemm.log 2>&1
echo "ncu rc=$?"; tail -7 gpurun_out/r2_qg_kbench.log
