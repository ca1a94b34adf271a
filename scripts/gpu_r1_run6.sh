set -x
timeout 900 python -m pytest tests/test_dynamic_gpu.py tests/test_sampler_gpu.py -x -q -s 2>&1 | grep -v Warning | grep "err\|passed\|failed\|FAILED\|Error\|error\|assert" | head -40 > gpurun_out/r1f_tests.log
timeout 600 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --dump-launches gpurun_out/r1f_launch_events.csv > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err
