#!/bin/bash
# one GPU: the whole GPU suite, then the default bench line
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/final_gpu_tests.log
python bench.py > gpurun_out/bench_final_1gpu.json 2> gpurun_out/bench_final_1gpu.err
cat gpurun_out/final_gpu_tests.log; tail -c 600 gpurun_out/bench_final_1gpu.json
