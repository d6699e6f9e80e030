#!/bin/bash
# bf16 epilogues with 64-column chunks: parity tests of the layered bf16 path, then the L step with and without
timeout 400 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_tc_gemm.py -x -q 2>&1 | tail -3
TSB200_OPT_OVERLAP=0 timeout 200 python scratch/lprof.py 2>&1 | grep -E "L model|gemm"
TSB200_GEMM_WIDE_EPI=0 TSB200_OPT_OVERLAP=0 timeout 200 python scratch/lprof.py 2>&1 | grep -E "L model|gemm"
timeout 200 python scratch/lprof.py 2>&1 | grep -E "L model"
