#!/bin/bash
# A/B of an environment knob inside one box: scripts/gpu_ab.sh NAME v1 v2 ...   (alternating runs, per-mode totals)
mkdir -p gpurun_out
K=$1; shift
for v in "$@" "$@"; do
  env $K=$v timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --dump-launches gpurun_out/ab_${K}_$v.csv 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$K=$v', round(d['value'],1), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()})"
done
