#!/bin/bash
# 3-D loop test with prints, then the whole GPU suite without -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_sampler_gpu.py -k "3d" -s -q > gpurun_out/r1n_3d.log 2>&1
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/r1n_tests.log 2>&1
tail -5 gpurun_out/r1n_tests.log
grep -E "^3d|FAILED|passed|failed" gpurun_out/r1n_3d.log
