#!/bin/bash
# Round-2 run C: tests touched since run B, the bench with its sections, warp-layout A/B, then the ncu evidence pass.
mkdir -p gpurun_out
O=gpurun_out/r2c
timeout 900 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "Warning\|warnings.warn\|^$" | grep "max\|median\|err\|passed\|failed\|FAILED\|verdicts\|full runs\|Error\|assert\|beyond\|dense" | cut -c1-400 > ${O}_tests.log
tail -6 ${O}_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc $?"; tail -3 ${O}_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2c_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"], d["roofline"]["class_ms"])
    print("field", {k: (round(v["ms"], 1), round(v["frac"], 3), v["sm_mhz_during_kernel"]) for k, v in d["field"].items() if isinstance(v, dict)})
    print("maze3d", d["maze3d"]["trajectories_per_s"], d["maze3d"]["roofline"]["frac"], "sweep", d["sweep"]["trajectories_per_s"], d["sweep"]["nn_query_ms"], d["sweep"]["nn_query_share"])
    print("eager ref", {k: v.get("trajectories_per_s") for k, v in d["reference_gpu_eager"].items() if isinstance(v, dict)}, d["reference_gpu_eager"].get("speedup_same_config_cfg_b1024"))
except Exception as e:
    print("bench parse failed", e)
PY
W=$PWD/ramp_b200/libramp_b200_wg0.so
for cfg in "default 2" "$W 2" "default 1" "default 2" "$W 2" "default 1"; do
  set -- $cfg
  if [ $1 = default ]; then unset RAMP_B200_LIB; else export RAMP_B200_LIB=$1; fi
  RAMP_UNET_SPLIT=$2 timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-sections --eager 2>> ${O}_ab.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1'.split('/')[-1], 'split $2', 'traj/s', round(d['value'],1), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_ab.log
done
unset RAMP_B200_LIB
bash scripts/gpu_r2_ncu.sh 2>&1 | tail -25
