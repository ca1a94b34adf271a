#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -x -q 2>&1 | tail -5 > gpurun_out/r1q_tests.log
cat gpurun_out/r1q_tests.log
grep -q failed gpurun_out/r1q_tests.log && exit 1
timeout 400 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --replan-ticks 10 --dump-launches gpurun_out/r1q_events.csv > gpurun_out/r1q_bench.json 2> gpurun_out/r1q_bench.err
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -4 >> gpurun_out/r1q_tests.log
tail -4 gpurun_out/r1q_tests.log
