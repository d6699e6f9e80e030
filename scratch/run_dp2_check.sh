#!/bin/bash
# two GPUs: the data-parallel GPU tests, then config 4's per-GPU share (8,192 rows) under the exchange schedules
timeout 900 python -m pytest tests/test_gpu_dp.py -x -q 2>&1 | tail -4
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $T --nproc-per-node 2 --master-port 29611 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
TSB200_NVLS_OVERLAP_SMS=0 timeout 200 $T --nproc-per-node 2 --master-port 29612 scratch/ldp.py nvls 8192 2>&1 | grep "L model"
timeout 200 $T --nproc-per-node 2 --master-port 29613 scratch/ldp.py p2p 8192 2>&1 | grep "L model"
