#!/bin/bash
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $T --nproc-per-node 1 --master-port 29681 scratch/dp_handler_probe.py 2>&1 | grep -E "us/batch|per epoch|Error|error"
timeout 300 $T --nproc-per-node 2 --master-port 29682 scratch/dp_handler_probe.py 2>&1 | grep -E "us/batch|per epoch|Error|error"
