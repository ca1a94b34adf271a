#!/bin/bash
# Round-2 run T: 32-wide k-blocks (SWIZZLE_64B operand tiles) for contractions with K <= 32 (RAMP_GEMM_BK32=1): parity tests,
# knob bit-identity, per-tag times and tick / headline A/B.
mkdir -p gpurun_out
O=gpurun_out/r2t
RAMP_GEMM_BK32=1 timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_fullsize_gpu.py tests/test_dynamic_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_tests.log
grep -q "failed\|error" ${O}_tests.log && exit 1
for v in 1 0; do echo "== RAMP_GEMM_BK32=$v"; RAMP_GEMM_BK32=$v timeout 300 python scripts/time_unet_tags.py '_K(16|32)x' 2>&1 | tail -16; done | tee ${O}_tags.txt
for v in 1 0 1 0; do
  echo "== RAMP_GEMM_BK32=$v"; RAMP_GEMM_BK32=$v timeout 300 python scripts/profile_tick.py 2>/dev/null | grep "per tick\|per level"
done | tee ${O}_tick.txt
for v in 1 0 1 0; do
  RAMP_GEMM_BK32=$v timeout 600 python bench.py --steps 3 --warmup 3 --no-sections 2> ${O}_bk_$v.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('RAMP_GEMM_BK32=$v', 'traj/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],1), d['clocks']['sm_mhz'])
"
done 2>&1 | tee ${O}_ab.txt
