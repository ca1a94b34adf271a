set -x
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -x -q -s 2>&1 | tail -30 > gpurun_out/r1c_tests.log
timeout 600 python -m pytest tests/test_sampler_gpu.py -x -q -s 2>&1 | grep -v Warning | tail -30 > gpurun_out/r1c_tests2.log
timeout 600 python bench.py --steps 2 --warmup 3 --cpu-sample 0 > gpurun_out/r1c_bench.json 2> gpurun_out/r1c_bench.err
