#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r1y
cap() { # name regex skip
  timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:$2 --launch-skip $3 --launch-count 1 -f -o ${O}_$1 python scripts/profile_unet.py > ${O}_ncu_$1.log 2>&1
}
cap gnb_c32 gn_backward 47
cap gnb_c256 gn_backward 63
cap gnf_c32 gn_forward 41
cap attnb64 attn_bwd_tc 32
cap attnf64 attn_fwd_tc 16
ls -la ${O}_*.ncu-rep
