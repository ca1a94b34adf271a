set -x
timeout 900 python -m pytest tests/test_unet_gpu.py -x -q -s 2>&1 | tail -6 > gpurun_out/r1i_tests.log
timeout 900 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err
