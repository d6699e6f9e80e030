#!/bin/bash
# CTA-pair GEMM: correctness first (under timeout: a barrier mistake would hang), then timings against one CTA per tile
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_tc_gemm.py -x -q 2>&1 | tail -5
echo "--- pair=1"; timeout 120 python scratch/gemm_pair.py
echo "--- pair=0"; TSB200_GEMM_PAIR=0 timeout 120 python scratch/gemm_pair.py
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "--- L step"; 
timeout 200 $T --nproc-per-node 1 --master-port 29601 scratch/ldp.py p2p 8192 2>&1 | grep "L model"
TSB200_GEMM_PAIR=0 timeout 200 $T --nproc-per-node 1 --master-port 29602 scratch/ldp.py p2p 8192 2>&1 | grep "L model"
