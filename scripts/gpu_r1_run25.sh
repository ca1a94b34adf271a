#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fullsize_gpu.py -x -q -s 2>&1 | grep -v Warning | tail -25 > gpurun_out/r1_fullsize.log
cat gpurun_out/r1_fullsize.log
