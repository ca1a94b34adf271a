#!/bin/bash
# Round-2 run O: per-launch staging heuristic (default) and the GEGLU forward on the 8-warp / 32-column epilogue
# configuration (RAMP_GEMM_E1_CW=32): parity tests with it, per-tag times, headline A/B.
mkdir -p gpurun_out
O=gpurun_out/r2o
RAMP_GEMM_E1_CW=32 timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_tests.log
grep -q "failed\|error" ${O}_tests.log && exit 1
run() { echo "== $1"; env $2 timeout 300 python scripts/time_unet_tags.py 'L(64|32)_e[1234]' 2>&1 | tail -19; }
{ run default "X=1"; run e1cw32 "RAMP_GEMM_E1_CW=32"; run e1cw32_2buf "RAMP_GEMM_E1_CW=32 RAMP_GEMM_STG_BUFS=1,2,2,2,2"; run allone "RAMP_GEMM_STG_BUFS=1,1,1,1,1"; } > ${O}_tags.txt 2>&1
grep "==\|per call\|e1" ${O}_tags.txt
for v in default e1cw32 default e1cw32; do
  e="X=1"; [ $v = e1cw32 ] && e="RAMP_GEMM_E1_CW=32"
  env $e timeout 600 python bench.py --steps 3 --warmup 3 --no-sections > ${O}_bench_$v.json 2> ${O}_bench_$v.err
  python - $v <<'PY'
import json, sys
try:
    d = json.loads(open("gpurun_out/r2o_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"], d["roofline"]["class_ms"])
except Exception as e:
    print("bench parse failed", e)
PY
done 2>&1 | tee ${O}_ab.txt
