#!/bin/bash
# Final round-1 evidence run: GPU test suite, both bench arms, smoke, ncu launch list of the bench command.
mkdir -p gpurun_out
O=gpurun_out/r1h
timeout 300 python -m pytest tests/test_pursuit.py tests/test_verdicts.py -m gpu -q -x -s 2>&1 | grep -v Warning | tail -25 > ${O}_tests_new.log
timeout 700 python -m pytest tests -m gpu -q 2>&1 | tail -6 > ${O}_tests.log
timeout 600 python bench.py > ${O}_bench.json 2> ${O}_bench.err
timeout 600 python bench.py --impl reference > ${O}_bench_reference.json 2> ${O}_bench_reference.err
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > ${O}_smoke.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file ${O}_launches.csv \
  python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field > ${O}_ncu_bench.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"pursuer_tick|box_points|nn_pack|apf_|cost_terms|select_best|smooth_segment|q_sample" -c 200 --csv --log-file ${O}_tick_launches.csv \
  python -m pytest tests/test_pursuit.py -m gpu -q -k "replan_graph or box_kernel" > ${O}_ncu_tick.log 2>&1
cat ${O}_tests_new.log; cat ${O}_tests.log; tail -2 ${O}_smoke.log
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r1h_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "replan", d.get("replan", {}).get("p50_ms"), d["clocks"])
PY
head -c 300 ${O}_bench_reference.json
