set -x
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_sampler_gpu.py -x -q 2>&1 | tail -4 > gpurun_out/r1e_tests.log
timeout 600 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --dump-launches gpurun_out/r1e_launch_events.csv > gpurun_out/r1e_bench.json 2> gpurun_out/r1e_bench.err
