#!/bin/bash
mkdir -p gpurun_out
timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_scene_encoder_gpu.py -x -q -s -k "oracle and 1" 2>&1 | grep -v Warning | tail -15 > gpurun_out/r1_scene_sani.log
tail -8 gpurun_out/r1_scene_sani.log
timeout 600 python -m pytest tests/test_scene_encoder_gpu.py -q -s 2>&1 | grep -v Warning | tail -25 > gpurun_out/r1_scene.log
cat gpurun_out/r1_scene.log
