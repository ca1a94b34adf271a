#!/bin/bash
# Round-2 run I: which side binds the HBM-bound GEMM epilogues?  Per-tag launch times of one 7168-row network evaluation with
# the product library and with the diagnostic builds (RAMP_GEMM_DIAG: 1 = no epilogue global loads, 2 = no epilogue global
# stores, 3 = neither; results wrong, timing only).  Plus the nearest-point kernel with its new default choice.
mkdir -p gpurun_out
O=gpurun_out/r2i
for v in default diag1 diag2 diag3; do
  lib=""; [ $v != default ] && lib=$PWD/ramp_b200/libramp_b200_$v.so
  echo "== $v"
  RAMP_B200_LIB=$lib timeout 300 python scripts/time_unet_tags.py 'L(64|8)_e[01234]' 2>&1 | tail -40
done > ${O}_diag.txt 2>&1
cat ${O}_diag.txt
for d in 2 3; do timeout 200 python scripts/time_field.py $d; done 2>&1 | tee ${O}_nn.txt
