#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dynamic_gpu.py -q -x -k streams 2>&1 | grep -v "Warning\|reg(" | head -60 > gpurun_out/r1_dyn.log
cat gpurun_out/r1_dyn.log
