#!/bin/bash
# Round-2 final evidence run (TMA-store epilogues, rotating staging tiles, relaxed pair arrive, FFMA2 nearest-point scan,
# attention backward with aliased P / dS): whole GPU suite, nearest-point timings, the bench line with every section, the
# reference arm, smoke, a same-box A/B of the staging heuristic, then the ncu evidence pass.
mkdir -p gpurun_out
O=gpurun_out/r2z
timeout 1200 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "Warning\|warnings.warn\|^$" | grep "max\|median\|err\|passed\|failed\|FAILED\|verdicts\|full runs\|Error\|assert\|beyond\|dense" | cut -c1-400 > ${O}_tests.log
tail -6 ${O}_tests.log
{ for dim in 2 3; do timeout 120 python scripts/time_field.py $dim; done; } > ${O}_nn.txt 2>&1; cat ${O}_nn.txt
timeout 1200 python bench.py --steps 5 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc $?"; tail -3 ${O}_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2z_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"], d["roofline"]["class_ms"])
    print("field", {k: (round(v["ms"], 1), round(v["frac"], 3), v["sm_mhz_during_kernel"]) for k, v in d["field"].items() if isinstance(v, dict)})
    print("replan", d["replan"]["p50_ms"], "replan25", d["replan25"]["p50_ms"], "maze3d", d["maze3d"]["trajectories_per_s"], d["maze3d"]["roofline"]["frac"], "sweep", d["sweep"]["trajectories_per_s"], d["sweep"]["nn_query_ms"], d["sweep"]["nn_query_share"])
    print("eager ref", {k: v.get("trajectories_per_s") for k, v in d["reference_gpu_eager"].items() if isinstance(v, dict)}, d["reference_gpu_eager"].get("speedup_same_config_cfg_b1024"))
except Exception as e:
    print("bench parse failed", e)
PY
timeout 900 python bench.py --impl reference --steps 5 --warmup 2 > ${O}_bench_reference.json 2> ${O}_bench_reference.err; echo "ref rc $?"; head -c 300 ${O}_bench_reference.json; echo
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for v in default one default one; do
  e="X=1"; [ $v = one ] && e="RAMP_GEMM_STG_BUFS=1,1,1,1,1"
  env $e timeout 600 python bench.py --steps 3 --warmup 3 --no-sections > ${O}_ab_$v.json 2> ${O}_ab_$v.err
  python - $v <<'PY'
import json, sys
try:
    d = json.loads(open("gpurun_out/r2z_ab_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
    print("staging", sys.argv[1], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"], d["roofline"]["class_ms"])
except Exception as e:
    print("bench parse failed", e)
PY
done 2>&1 | tee ${O}_stg_ab.txt
O=gpurun_out/r2zn
NCU="ncu --clock-control none"
for d in 2 3; do
  python scripts/profile_field.py $d > ${O}_field$d.plain 2>&1 && \
  $NCU --set full --import-source on -k regex:nn_query_kernel -s 2 -c 1 -f -o ${O}_nn$d python scripts/profile_field.py $d > ${O}_nn$d.log 2>&1
  echo "nn d=$d rc $?"
  ncu -i ${O}_nn$d.ncu-rep --page raw --csv > ${O}_nn$d.raw.csv 2>/dev/null
  rm -f ${O}_nn$d.ncu-rep
done
export RAMP_UNET_SPLIT=1       # whole 7168-row calls (one chain of 260 GEMM launches) for the per-launch evidence
python scripts/profile_unet.py > ${O}_unet.plain 2>&1 && \
$NCU --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --kernel-name regex:gemm_tc_kernel \
    --launch-skip 260 --launch-count 260 --csv --log-file ${O}_gemm_traffic.csv python scripts/profile_unet.py > ${O}_traffic.log 2>&1
echo "traffic rc $?"
full() {  # name kernel-regex skip
  $NCU --set full --kernel-name-base demangled -k "regex:$2" -s $3 -c 1 -f -o ${O}_$1 python scripts/profile_unet.py > ${O}_$1.log 2>&1
  echo "$1 rc $?"
  ncu -i ${O}_$1.ncu-rep --page raw --csv > ${O}_$1.raw.csv 2>/dev/null
  rm -f ${O}_$1.ncu-rep      # gpurun_out/ is limited to 64 MiB: keep the raw page
}
full e1 "gemm_tc_kernel<.int.1," 16
full e2 "gemm_tc_kernel<.int.2," 30
full e4ff "gemm_tc_kernel<.int.4," 60
full attn_bwd64 "attn_bwd_tc<.int.64," 3
unset RAMP_UNET_SPLIT
python bench.py --steps 1 --warmup 3 --cpu-sample 0 --no-sections --eager > ${O}_bench.plain 2> ${O}_bench.plain.err && \
$NCU --metrics gpu__time_duration.sum --launch-skip 7700 --launch-count 7600 --csv --log-file ${O}_launches.csv \
    python bench.py --steps 1 --warmup 3 --cpu-sample 0 --no-sections --eager > ${O}_ncu_bench.log 2>&1
echo "launch list rc $?"
ls -la gpurun_out | grep r2z
