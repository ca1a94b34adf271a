#!/bin/bash
# Round-2 run H: nearest-point scan on packed FP32 FMAs (FFMA2, point pairs in shared memory) — exactness tests and the
# (R, TILE) sweep; attention backward with P / dS parked over K / V (3 CTAs per SM) — network parity tests and a same-box A/B
# of the attention class time against the library without the aliasing.
mkdir -p gpurun_out
O=gpurun_out/r2h
timeout 600 python -m pytest tests/test_loops_gpu.py tests/test_sampler_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x -k "nn_query or nearest or cloud or field" 2>&1 | tail -4 | tee ${O}_nn_tests.log
timeout 900 python scripts/sweep_field.py > ${O}_nn_sweep.txt 2>&1; cat ${O}_nn_sweep.txt
timeout 600 python -m pytest tests/test_unet_gpu.py tests/test_gemm_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_unet_tests.log
for v in alias noalias alias noalias; do
  lib=""; [ $v = noalias ] && lib=$PWD/ramp_b200/libramp_b200_noalias.so
  RAMP_B200_LIB=$lib timeout 600 python bench.py --steps 3 --warmup 3 --no-sections > ${O}_bench_$v.json 2> ${O}_bench_$v.err
  python - $v <<'PY'
import json, sys
try:
    d = json.loads(open("gpurun_out/r2h_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"], d["roofline"]["class_ms"])
except Exception as e:
    print("bench parse failed", e)
PY
done 2>&1 | tee ${O}_attn_ab.txt
