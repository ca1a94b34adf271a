set -x
timeout 900 python -m pytest tests/test_unet_gpu.py tests/test_gemm_gpu.py -x -q 2>&1 | tail -3 > gpurun_out/r1g_tests.log
timeout 900 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 40 --dump-launches gpurun_out/r1g_launch_events.csv > gpurun_out/r1g_bench.json 2> gpurun_out/r1g_bench.err
