#!/bin/bash
# Evidence run after the f4 work (device pursuer, duplicate-row collapse of the dynamic model): GPU test suite, both bench
# arms, smoke, and the ncu launch list of the bench command.
mkdir -p gpurun_out
O=gpurun_out/r1g
timeout 300 python -m pytest tests/test_pursuit.py tests/test_dynamic_gpu.py -m gpu -q -x -s 2>&1 | tail -15 > ${O}_tests_new.log
timeout 700 python -m pytest tests -m gpu -q 2>&1 | tail -6 > ${O}_tests.log
timeout 600 python bench.py > ${O}_bench.json 2> ${O}_bench.err
timeout 600 python bench.py --impl reference > ${O}_bench_reference.json 2> ${O}_bench_reference.err
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > ${O}_smoke.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file ${O}_launches.csv \
  python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field > ${O}_ncu_bench.log 2>&1
cat ${O}_tests_new.log; cat ${O}_tests.log; tail -2 ${O}_smoke.log; head -c 300 ${O}_bench.json; echo
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r1g_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "replan", d.get("replan", {}).get("p50_ms"), d.get("replan", {}).get("network_rows_per_step"))
PY
head -c 300 ${O}_bench_reference.json
