#!/bin/bash
mkdir -p gpurun_out
{ for a in "2 8192" "3 8192" "2 1024" "3 1024" "2 64" "2 35"; do timeout 120 python scripts/time_field.py $a; done; } > gpurun_out/nn_auto.txt 2>&1
cat gpurun_out/nn_auto.txt
timeout 600 python -m pytest tests/test_sampler_gpu.py tests/test_fullsize_gpu.py tests/test_dropin_gpu.py tests/test_dynamic_gpu.py -q 2>&1 | tail -2
