#!/bin/bash
# Round-2 run G: nearest-point kernel with its chunk bookkeeping in shared memory — exactness tests, the (R, TILE) sweep at the
# sweep's shape, and the per-launch CSV of one 7168-row headline step (shapes + CUDA-event times) for the GEMM floor model.
mkdir -p gpurun_out
O=gpurun_out/r2g
timeout 600 python -m pytest tests/test_loops_gpu.py tests/test_sampler_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x -k "nn_query or nearest or cloud or field" 2>&1 | tail -4 | tee ${O}_nn_tests.log
timeout 900 python scripts/sweep_field.py > ${O}_nn_sweep.txt 2>&1; cat ${O}_nn_sweep.txt
timeout 600 python bench.py --steps 2 --warmup 3 --no-sections --dump-launches ${O}_launches.csv > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc $?"; tail -2 ${O}_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2g_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"], d["roofline"]["class_ms"])
except Exception as e:
    print("bench parse failed", e)
PY
wc -l ${O}_launches.csv; head -3 ${O}_launches.csv
