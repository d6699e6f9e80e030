#!/bin/bash
# usage: run_ncu_l.sh list|full   (one ncu pass per gpurun call) -- two fits of the L model at 8,192 rows
if [ "$1" = list ]; then
  timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2p_L_launches.csv python scratch/l2fits.py > gpurun_out/r2p_ncu_list.log 2>&1
else
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:'tc_gemm_kernel|head_dh' --launch-skip 16 --launch-count 12 -o gpurun_out/r2p_L_full -f python scratch/l2fits.py > gpurun_out/r2p_ncu_full.log 2>&1
fi
tail -2 gpurun_out/r2p_ncu_*.log
