#!/bin/bash
# Round-2 run M: TMA store path of the GEMM epilogue: parity tests, then same-box A/B (RAMP_GEMM_TMA_STORE=1 / 0) of the
# headline and the per-tag launch times at L = 64.
mkdir -p gpurun_out
O=gpurun_out/r2m
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_unet_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x 2>&1 | tail -15 | tee ${O}_tests.log
grep -q "failed\|error" ${O}_tests.log && exit 1
for v in 1 0 1 0; do
  RAMP_GEMM_TMA_STORE=$v timeout 600 python bench.py --steps 3 --warmup 3 --no-sections > ${O}_bench_$v.json 2> ${O}_bench_$v.err
  python - $v <<'PY'
import json, sys
try:
    d = json.loads(open("gpurun_out/r2m_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print("tma_store", sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"], d["roofline"]["class_ms"])
except Exception as e:
    print("bench parse failed", e)
PY
done 2>&1 | tee ${O}_ab.txt
for v in 1 0; do echo "== RAMP_GEMM_TMA_STORE=$v"; RAMP_GEMM_TMA_STORE=$v timeout 300 python scripts/time_unet_tags.py 'L(64|8)_e[01234]' 2>&1 | tail -45; done > ${O}_tags.txt 2>&1
