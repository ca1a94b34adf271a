#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/profile_unet.py > gpurun_out/r1p_plain.log 2>&1 || exit 1
for spec in e1:268 e2:506 e4ff:507 e3:267; do
  name=${spec%%:*}; skip=${spec##*:}
  timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:gemm_tc_kernel \
    --launch-skip $skip --launch-count 1 -f -o gpurun_out/r1p_gemm_$name python scripts/profile_unet.py > gpurun_out/r1p_ncu_$name.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
