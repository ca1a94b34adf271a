set -x
timeout 600 python -m pytest tests/test_sampler_gpu.py -x -q -s 2>&1 | grep -v Warning | grep "free-running\|passed\|failed\|FAILED\|teacher" > gpurun_out/r1d_tests2.log
timeout 600 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --dump-launches gpurun_out/r1d_launch_events.csv > gpurun_out/r1d_bench.json 2> gpurun_out/r1d_bench.err
