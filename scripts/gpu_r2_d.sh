#!/bin/bash
# Round-2 run D: CTA pairs (tcgen05 cta_group::2, RAMP_GEMM_PAIR=1) — parity, then same-box A/B of the headline.
mkdir -p gpurun_out
O=gpurun_out/r2d
timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x 2>&1 | tail -3 | tee ${O}_gemm_pair0.log
RAMP_GEMM_PAIR=1 timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x 2>&1 | tail -8 | tee ${O}_gemm_pair1.log
RAMP_GEMM_PAIR=1 timeout 600 python -m pytest tests/test_unet_gpu.py tests/test_fullsize_gpu.py tests/test_verdicts.py -m gpu -q -s 2>&1 | grep "err\|passed\|failed\|FAILED\|verdicts\|Error" | cut -c1-300 | tee ${O}_tests_pair1.log
for cfg in 0 1 0 1; do
  RAMP_GEMM_PAIR=$cfg timeout 300 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-sections --eager 2>> ${O}_ab.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('pair $cfg traj/s', round(d['value'],1), {k: round(x,1) for k,x in d['roofline']['class_ms'].items()}, d['clocks']['sm_mhz'])
" | tee -a ${O}_ab.log
done
RAMP_GEMM_PAIR=1 timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --no-sections --eager --dump-launches ${O}_launches_pair1.csv > /dev/null 2>&1
RAMP_GEMM_PAIR=0 timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --no-sections --eager --dump-launches ${O}_launches_pair0.csv > /dev/null 2>&1
python - <<'PY'
import csv, collections, re
for tag in ("pair0", "pair1"):
    agg = collections.defaultdict(lambda: [0, 0.0])
    try:
        for r in csv.DictReader(open("gpurun_out/r2d_launches_%s.csv" % tag)):
            if r["class"] != "gemm":
                continue
            e = re.search(r"_e(\d)", r["tag"]); L = re.search(r"_L(\d+)", r["tag"])
            agg[("e" + e.group(1), int(L.group(1)))][0] += 1
            agg[("e" + e.group(1), int(L.group(1)))][1] += float(r["ms"])
        print(tag, {k: round(v[1], 1) for k, v in sorted(agg.items())})
    except Exception as ex:
        print(tag, "failed", ex)
PY
# nearest-point kernel: tracked chunks (default = top 3 + value of the 4th, libramp_b200_nn2.so = top 2 + value of the 3rd), R sweep
{
N2=$PWD/ramp_b200/libramp_b200_nn2.so
for lib in default $N2; do
  if [ $lib = default ]; then unset RAMP_B200_LIB; else export RAMP_B200_LIB=$lib; fi
  for dim in 2 3; do
    echo -n "$(basename $lib) auto: "; timeout 120 python scripts/time_field.py $dim
    for r in 4 6 8 10 12; do
      echo -n "$(basename $lib) R=$r: "; RAMP_NN_FORCE=$r,1024 timeout 120 python scripts/time_field.py $dim
    done
  done
done
unset RAMP_B200_LIB
} > ${O}_nn_sweep.txt 2>&1
cat ${O}_nn_sweep.txt
timeout 300 python -m pytest tests/test_loops_gpu.py tests/test_sampler_gpu.py tests/test_fullsize_gpu.py -m gpu -q -k "nn_query" 2>&1 | tail -2
