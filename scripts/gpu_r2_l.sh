#!/bin/bash
# Round-2 run L: relaxed remote arrive + batched staging read-backs in the GEMM epilogue: GEMM / network tests, then a
# same-box A/B of the headline against the library with the .release arrive (both have the batched read-backs).
mkdir -p gpurun_out
O=gpurun_out/r2l
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_tests.log
for v in new release new release; do
  lib=""; [ $v = release ] && lib=$PWD/ramp_b200/libramp_b200_relarr.so
  RAMP_B200_LIB=$lib timeout 600 python bench.py --steps 3 --warmup 3 --no-sections > ${O}_bench_$v.json 2> ${O}_bench_$v.err
  python - $v <<'PY'
import json, sys
try:
    d = json.loads(open("gpurun_out/r2l_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"], d["roofline"]["class_ms"])
except Exception as e:
    print("bench parse failed", e)
PY
done 2>&1 | tee ${O}_ab.txt
timeout 300 python scripts/time_unet_tags.py 'L64_e[01234]' 2>&1 | tail -22 | tee ${O}_tags.txt
