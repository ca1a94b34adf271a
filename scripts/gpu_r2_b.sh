#!/bin/bash
# Round-2 run B: full GPU test suite with the fp16 + fp16 planes (satfinite split), the new bench sections, the
# two-stream experiment, the reference arm.
mkdir -p gpurun_out
O=gpurun_out/r2b
timeout 900 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "Warning\|warnings.warn\|^$" | grep "max\|median\|err\|passed\|failed\|FAILED\|verdicts\|full runs\|Error\|assert\|beyond\|dense" | cut -c1-700 > ${O}_tests.log
tail -12 ${O}_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc $?"; tail -3 ${O}_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2b_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"])
    for k in ("replan", "replan25", "maze3d", "sweep", "field", "reference_gpu_eager", "scene_encoder", "cpu_baseline"):
        print(k, json.dumps(d.get(k))[:900])
except Exception as e:
    print("bench parse failed", e)
PY
timeout 600 python scripts/exp_two_streams.py 2 2>&1 | tail -2 | tee ${O}_two_streams.log
timeout 600 python scripts/exp_two_streams.py 4 2>&1 | tail -1 | tee -a ${O}_two_streams.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > ${O}_bench_reference.json 2> ${O}_bench_reference.err; echo "ref rc $?"
head -c 1500 ${O}_bench_reference.json
# engine-level row split (RAMP_UNET_SPLIT): same-box A/B of the headline, eager loop
for sp in 1 2 1 2 3; do
  RAMP_UNET_SPLIT=$sp timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-sections 2>> ${O}_split.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('split $sp traj/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_split.log
done
RAMP_UNET_SPLIT=2 timeout 300 python -m pytest tests/test_unet_gpu.py tests/test_fullsize_gpu.py -m gpu -q 2>&1 | tail -3 | tee -a ${O}_split.log
# A/B: libramp_b200_base.so (HEAD at 6f0f... = before) vs the default library (two epilogue loads in flight in the GEGLU- /
# LayerNorm-backward epilogues, __expf in the GEGLU forward, bounded mbarrier waits)
BASE=$PWD/ramp_b200/libramp_b200_base.so
MID=$PWD/ramp_b200/libramp_b200_mid.so      # two loads in flight, round-1 warp layout (spills); default = + warp-group roles with setmaxnreg
for lib in $BASE $MID default $BASE $MID default; do
  if [ $lib = default ]; then unset RAMP_B200_LIB; else export RAMP_B200_LIB=$lib; fi
  timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-sections 2>> ${O}_epi.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib'.split('/')[-1], 'traj/s', round(d['value'],1), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_epi.log
done
unset RAMP_B200_LIB
timeout 300 python scripts/profile_tick.py > ${O}_tick_levels.txt 2> ${O}_tick_levels.err; tail -4 ${O}_tick_levels.txt
