set -x
timeout 240 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -x -q 2>&1 | tail -3 > gpurun_out/r1j_tests.log
timeout 420 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --dump-launches gpurun_out/r1j_launch_events.csv > gpurun_out/r1j_bench.json 2> gpurun_out/r1j_bench.err
