#!/bin/bash
# L model, 8,192 rows per GPU (the per-GPU share of config 4 at 8 GPUs), two GPUs: every exchange schedule next to one GPU alone
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$T --nproc-per-node 1 --master-port 29601 scratch/ldp.py p2p 8192 2>&1 | grep "L model"
$T --nproc-per-node 2 --master-port 29602 scratch/ldp.py nccl 8192 2>&1 | grep "L model"
TSB200_NVLS_OVERLAP_SMS=0 $T --nproc-per-node 2 --master-port 29603 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
TSB200_NVLS_OVERLAP_SMS=8 $T --nproc-per-node 2 --master-port 29604 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
TSB200_NVLS_OVERLAP_SMS=16 $T --nproc-per-node 2 --master-port 29605 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
TSB200_NVLS_OVERLAP_SMS=24 $T --nproc-per-node 2 --master-port 29606 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
