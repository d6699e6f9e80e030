#!/bin/bash
# config 4 on one GPU (65,536 rows per step): CTA-pair GEMM against one CTA per tile, twice each, alternating
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for i in 1 2; do
timeout 200 $T --nproc-per-node 1 --master-port 2961$i scratch/ldp.py p2p 65536 2>&1 | grep "L model"
TSB200_GEMM_PAIR=0 timeout 200 $T --nproc-per-node 1 --master-port 2962$i scratch/ldp.py p2p 65536 2>&1 | grep "L model"
done
