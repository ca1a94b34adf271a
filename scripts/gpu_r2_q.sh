#!/bin/bash
# Round-2 run Q: wave-aligned row split (RAMP_UNET_SPLIT_ALIGN=1: 7168 rows as 4704 + 2352 + 112 instead of 3584 + 3584):
# parity tests under it, then same-box A/B of the headline (equal split / aligned / no split).
mkdir -p gpurun_out
O=gpurun_out/r2q
RAMP_UNET_SPLIT_ALIGN=1 timeout 900 python -m pytest tests/test_fullsize_gpu.py tests/test_loops_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_tests.log
grep -q "failed\|error" ${O}_tests.log && exit 1
for v in equal aligned equal aligned nosplit aligned3; do
  e="X=1"; [ $v = aligned ] && e="RAMP_UNET_SPLIT_ALIGN=1"; [ $v = nosplit ] && e="RAMP_UNET_SPLIT=1"; [ $v = aligned3 ] && e="RAMP_UNET_SPLIT_ALIGN=1 RAMP_UNET_SPLIT=3"
  env $e timeout 600 python bench.py --steps 3 --warmup 3 --no-sections > ${O}_bench_$v.json 2> ${O}_bench_$v.err
  python - $v <<'PY'
import json, sys
try:
    d = json.loads(open("gpurun_out/r2q_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print("split", sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms/step", round(d["ms_per_step"], 1), d["clocks"]["sm_mhz"])
except Exception as e:
    print("bench parse failed", e)
PY
done 2>&1 | tee ${O}_ab.txt
