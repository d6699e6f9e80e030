#!/bin/bash
# Config 4 on 8 GPUs, 8,192 rows per GPU: exchange schedules the bench does not cover (it times nvls / nvls_overlap(20) / nccl itself)
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
TSB200_NVLS_OVERLAP_SMS=12 $T --master-port 29701 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
TSB200_NVLS_OVERLAP_SMS=28 $T --master-port 29702 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
