#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dynamic_gpu.py -q -x 2>&1 | grep -v Warning | tail -12 > gpurun_out/r1_dyn.log
cat gpurun_out/r1_dyn.log
timeout 400 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 30 --no-field > gpurun_out/r1_bench_t.json 2> gpurun_out/r1_bench_t.err
python -c "
import json
d=json.load(open('gpurun_out/r1_bench_t.json'))
print(d['value']); print(d['replan'])" ; tail -3 gpurun_out/r1_bench_t.err
