set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r1_tests.log
python bench.py --steps 3 --warmup 3 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r1_ref.json 2> gpurun_out/r1_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 1 --warmup 3 --cpu-sample 0 > gpurun_out/r1_ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc --launch-skip 300 -c 12 -o gpurun_out/r1_gemm_full -f python bench.py --steps 1 --warmup 3 --cpu-sample 0 > gpurun_out/r1_ncu_full.log 2>&1
ls -la gpurun_out
