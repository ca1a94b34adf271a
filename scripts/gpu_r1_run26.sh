#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r1y
for spec in conv5n32:261 qkv:266 e3:267; do
  name=${spec%%:*}; skip=${spec##*:}
  timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:gemm_tc_kernel \
    --launch-skip $skip --launch-count 1 -f -o ${O}_gemm_$name python scripts/profile_unet.py > ${O}_ncu_$name.log 2>&1
done
ls -la ${O}_gemm_*.ncu-rep
