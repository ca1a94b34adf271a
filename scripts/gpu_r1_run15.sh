#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_unet_gpu.py -s -q > gpurun_out/r1o_unet.log 2>&1
for i in 1 2 3; do timeout 200 python -m pytest tests/test_sampler_gpu.py -k "3d_ddpm" -s -q 2>&1 | grep -E "^3d free|passed|failed" ; done > gpurun_out/r1o_3d.log
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/r1o_tests.log
timeout 400 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --replan-ticks 10 > gpurun_out/r1o_bench.json 2> gpurun_out/r1o_bench.err
grep -E "err|passed|failed" gpurun_out/r1o_unet.log; cat gpurun_out/r1o_3d.log gpurun_out/r1o_tests.log
