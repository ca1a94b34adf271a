#!/bin/bash
# Round-2 A/B of the operand-plane format inside one box: default library = fp16 + fp16 planes with a loss-scaled
# backward sweep, libramp_b200_bf16.so = round 1's bf16 + bf16 planes.  Build both first:
#     python -m ramp_b200.build --force && python -m ramp_b200.build --planes-bf16
mkdir -p gpurun_out
O=gpurun_out/r2a
timeout 900 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "Warning\|warnings.warn\|^$" | grep "max\|median\|err\|passed\|failed\|FAILED\|verdicts\|full runs\|Error\|assert" | cut -c1-600 > ${O}_tests_fp16.log
tail -40 ${O}_tests_fp16.log
B=$PWD/ramp_b200/libramp_b200_bf16.so
RAMP_B200_LIB=$B timeout 400 python -m pytest tests/test_verdicts.py tests/test_sampler_gpu.py tests/test_dynamic_gpu.py -m gpu -q -s 2>&1 | grep "max\|median\|passed\|failed\|FAILED\|verdicts\|full runs" | cut -c1-600 > ${O}_tests_bf16.log
tail -15 ${O}_tests_bf16.log
for lib in default $B default $B; do
  if [ $lib = default ]; then unset RAMP_B200_LIB; else export RAMP_B200_LIB=$lib; fi
  timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-field --replan-ticks 20 2> ${O}_bench.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib'.split('/')[-1], 'traj/s', round(d['value'],1), 'replan p50', round(d['replan']['p50_ms'],2), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_ab.log
done
