#!/bin/bash
# two GPUs, S-model data-parallel step: low-latency packets (one hop), the same with the owner hop, publish + remote reads; then the p2p tests
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for i in 1 2; do
timeout 200 $T --nproc-per-node 2 --master-port 2963$i scratch/kprof_dp.py p2p 2>&1 | grep -E "us per fit|CTA 0"
TSB200_DP_LL=2 timeout 200 $T --nproc-per-node 2 --master-port 2965$i scratch/kprof_dp.py p2p 2>&1 | grep -E "us per fit|CTA 0"
TSB200_DP_LL=0 timeout 200 $T --nproc-per-node 2 --master-port 2964$i scratch/kprof_dp.py p2p 2>&1 | grep -E "us per fit|CTA 0"
done
timeout 600 python -m pytest tests/test_gpu_dp.py -x -q -k "p2p or never_arrives" 2>&1 | tail -4
