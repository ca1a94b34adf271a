#!/bin/bash
# A/B of the experimental plane format (fp16 hi + bf16 residual, -DRAMP_HI_FP16=1) against the default library inside one
# box.  Build BOTH libraries in the build container first (they travel with the snapshot):
#     python -m ramp_b200.build --force && python -m ramp_b200.build --hi-fp16
# 1. GEMM / network-row / sampler / verdict parity with the variant (the free-run deviation line of test_verdicts is the number
#    the variant is meant to move: default 4.4e-4 median, 1.04e-3 max), 2. alternating bench runs.
mkdir -p gpurun_out
O=gpurun_out/hf16
V=$PWD/ramp_b200/libramp_b200_hf16.so
RAMP_B200_LIB=$V timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x 2>&1 | tail -5 > ${O}_gemm.log; cat ${O}_gemm.log
RAMP_B200_LIB=$V timeout 300 python -m pytest tests/test_unet_gpu.py tests/test_sampler_gpu.py -m gpu -q -s 2>&1 | grep -v "Warning\|reg(" | tail -25 > ${O}_rows.log; cat ${O}_rows.log
RAMP_B200_LIB=$V timeout 300 python -m pytest tests/test_verdicts.py -m gpu -q -s 2>&1 | grep "verdicts compared\|full runs\|passed\|failed" | cut -c1-400 > ${O}_verdicts.log; cat ${O}_verdicts.log
for lib in default $V default $V; do
  if [ $lib = default ]; then unset RAMP_B200_LIB; else export RAMP_B200_LIB=$lib; fi
  timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-field --replan-ticks 20 2> ${O}_bench.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib'.split('/')[-1], 'traj/s', round(d['value'],1), 'replan p50', round(d['replan']['p50_ms'],2), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_ab.log
done
