#!/bin/bash
# Round-2 run N: rotating staging tiles of the TMA store path (RAMP_GEMM_STG_BUFS = tiles per warp wanted by e0..e4):
# parity tests with the deepest rotation, then per-tag launch times of one 7168-row network call per configuration.
mkdir -p gpurun_out
O=gpurun_out/r2n2
RAMP_GEMM_STG_BUFS=3,3,3,3,3 timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_tests.log
grep -q "failed\|error" ${O}_tests.log && exit 1
for cfg in 1,1,1,1,1 1,2,2,2,2 1,3,2,2,1 2,3,3,3,3 2,2,2,2,2 1,3,3,3,2; do
  echo "== $cfg"
  RAMP_GEMM_STG_BUFS=$cfg timeout 300 python scripts/time_unet_tags.py 'L64_e[01234]' 2>&1 | tail -19
done > ${O}_tags.txt 2>&1
grep "==\|per call" ${O}_tags.txt
