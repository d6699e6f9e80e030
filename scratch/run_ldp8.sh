#!/bin/bash
# Config 4 on 8 GPUs, 8,192 rows per GPU: one whole-buffer nvls kernel vs the per-layer overlapped exchange (two shadow sets)
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
for sms in 0 20 16; do
  TSB200_NVLS_OVERLAP_SMS=$sms $T --master-port $((29700 + sms)) scratch/ldp.py nvls 8192 2>&1 | grep "L model"
done
