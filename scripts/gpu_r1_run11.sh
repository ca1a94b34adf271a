set -x
timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_dropin_gpu.py -x -q 2>&1 | tail -3 > gpurun_out/r1k_tests.log
timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --dump-launches gpurun_out/r1k_launch_events.csv > gpurun_out/r1k_bench.json 2> gpurun_out/r1k_bench.err
timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --max-rows 7168 > gpurun_out/r1k_bench_7168.json 2> gpurun_out/r1k_bench_7168.err
timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --max-rows 3584 > gpurun_out/r1k_bench_3584.json 2> gpurun_out/r1k_bench_3584.err
