#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -x -q 2>&1 | tail -3 > gpurun_out/r1u_tests.log
cat gpurun_out/r1u_tests.log
grep -q failed gpurun_out/r1u_tests.log && exit 1
timeout 400 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --dump-launches gpurun_out/r1u_events.csv > gpurun_out/r1u_bench.json 2> gpurun_out/r1u_bench.err
python -c "
import json
d=json.load(open('gpurun_out/r1u_bench.json'))
print(d['value'],d['ms_per_step'],d['roofline']['class_ms'])"
for spec in e1:268 e2:506 e4ff:507; do
  name=${spec%%:*}; skip=${spec##*:}
  timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:gemm_tc_kernel \
    --launch-skip $skip --launch-count 1 -f -o gpurun_out/r1u_gemm_$name python scripts/profile_unet.py > gpurun_out/r1u_ncu_$name.log 2>&1
done
