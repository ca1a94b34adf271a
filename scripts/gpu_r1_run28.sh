#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_scene_encoder_gpu.py -q -s 2>&1 | grep -v Warning | tail -25 > gpurun_out/r1_scene.log
cat gpurun_out/r1_scene.log
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > gpurun_out/r1_all.log
cat gpurun_out/r1_all.log
timeout 400 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 10 > gpurun_out/r1_bench_s.json 2> gpurun_out/r1_bench_s.err
python -c "
import json
d=json.load(open('gpurun_out/r1_bench_s.json'))
print(d['value'], d['scene_encoder']); print(d['replan']['p50_ms'])"
