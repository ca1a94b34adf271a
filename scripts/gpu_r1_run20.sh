#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -x -q 2>&1 | tail -5 > gpurun_out/r1w_tests.log
cat gpurun_out/r1w_tests.log
grep -q failed gpurun_out/r1w_tests.log && exit 1
timeout 400 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --dump-launches gpurun_out/r1w_events.csv > gpurun_out/r1w_bench.json 2> gpurun_out/r1w_bench.err
python -c "
import json
d=json.load(open('gpurun_out/r1w_bench.json'))
print(d['value'],d['ms_per_step'],d['roofline']['class_ms'])"
