#!/bin/bash
mkdir -p gpurun_out
for m in 0 1 2 4 8 16; do
  RAMP_GEMM_PREFETCH=$m timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --dump-launches gpurun_out/r1s_pf$m.csv 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$m', d['value'], d['roofline']['class_ms'])"
done > gpurun_out/r1s_variants.txt 2>&1
cat gpurun_out/r1s_variants.txt
