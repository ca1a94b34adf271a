#!/bin/bash
# Round-2 run R: where a replan tick's time goes on the final library (per level / class), and programmatic dependent launch
# once more after the epilogue work (same box, alternating).
mkdir -p gpurun_out
O=gpurun_out/r2r
timeout 300 python scripts/profile_tick.py > ${O}_tick_levels.txt 2> ${O}_tick.err; tail -8 ${O}_tick_levels.txt
for v in 0 7 0 7; do
  RAMP_PDL=$v timeout 600 python bench.py --steps 3 --warmup 3 --no-sections 2> ${O}_pdl_$v.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('RAMP_PDL=$v', 'traj/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],1), d['clocks']['sm_mhz'])
"
done 2>&1 | tee ${O}_pdl_ab.txt
