#!/bin/bash
# Round-2 closing run on the final library: whole GPU suite, bench line with every section, reference arm, smoke.
mkdir -p gpurun_out
O=gpurun_out/r2y
timeout 1200 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "Warning\|warnings.warn\|^$" | grep "max\|median\|err\|passed\|failed\|FAILED\|verdicts\|full runs\|Error\|assert\|beyond\|dense\|bit-identical" | cut -c1-400 > ${O}_tests.log
tail -4 ${O}_tests.log
timeout 1200 python bench.py --steps 5 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc $?"; tail -3 ${O}_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2y_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"], d["roofline"]["class_ms"])
    print("field", {k: (round(v["ms"], 1), round(v["frac"], 3), v["sm_mhz_during_kernel"]) for k, v in d["field"].items() if isinstance(v, dict)})
    print("replan", d["replan"]["p50_ms"], "replan25", d["replan25"]["p50_ms"], "maze3d", d["maze3d"]["trajectories_per_s"], d["maze3d"]["roofline"]["frac"], "sweep", d["sweep"]["trajectories_per_s"], d["sweep"]["nn_query_ms"], d["sweep"]["nn_query_share"])
    print("eager ref", {k: v.get("trajectories_per_s") for k, v in d["reference_gpu_eager"].items() if isinstance(v, dict)}, d["reference_gpu_eager"].get("speedup_same_config_cfg_b1024"))
except Exception as e:
    print("bench parse failed", e)
PY
timeout 900 python bench.py --impl reference --steps 5 --warmup 2 > ${O}_bench_reference.json 2> ${O}_bench_reference.err; echo "ref rc $?"; head -c 200 ${O}_bench_reference.json; echo
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
