#!/bin/bash
# A/B of programmatic dependent launch (RAMP_PDL bit mask: 1 GEMM, 2 GroupNorm, 4 attention) inside one box:
# parity tests with everything on, then alternating bench runs (sampling throughput + replan tick p50).
mkdir -p gpurun_out
O=gpurun_out/pdl
RAMP_PDL=7 timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_sampler_gpu.py tests/test_dynamic_gpu.py tests/test_fullsize_gpu.py tests/test_verdicts.py -m gpu -q -x 2>&1 | tail -4 > ${O}_tests.log
cat ${O}_tests.log
for v in 0 7 1 0 7 1; do
  RAMP_PDL=$v timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-field --replan-ticks 40 2> ${O}_$v.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('RAMP_PDL=$v', 'traj/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'replan p50', round(d['replan']['p50_ms'],2), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_ab.log
done
