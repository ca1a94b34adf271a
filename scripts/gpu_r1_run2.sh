set -x
python -m pytest tests/test_unet_gpu.py tests/test_sampler_gpu.py -x -q 2>&1 | tail -5 > gpurun_out/r1b_tests.log
python bench.py --steps 2 --warmup 3 --cpu-sample 0 > gpurun_out/r1b_bench.json 2> gpurun_out/r1b_bench.err
