#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_metrics_gpu.py -q -s 2>&1 | grep -v Warning | tail -20 > gpurun_out/r1_metrics.log
cat gpurun_out/r1_metrics.log
