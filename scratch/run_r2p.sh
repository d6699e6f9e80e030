#!/bin/bash
# after the pair GEMM + SIMT head dH: GPU suite, then the L step with its serialized kernel shares
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 200 python scratch/lprof.py
TSB200_HEAD_DH_GEMM=1 timeout 200 python scratch/lprof.py | head -1
