#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 600 python scratch/prof_mega.py > gpurun_out/r2_prof_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 4 -c 1 -f -o gpurun_out/r2_mega_v11 python scratch/prof_mega.py > gpurun_out/r2_ncu_mega.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/r2_prof_plain.log
