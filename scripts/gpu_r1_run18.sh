#!/bin/bash
# DRAM traffic of the GEGLU-backward (gemm launch 246) and FF1-backward (247) launches under tuning knobs
mkdir -p gpurun_out
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
run() { # name, env...
  local name=$1; shift
  env "$@" timeout 300 ncu --metrics $M --clock-control none --kernel-name regex:gemm_tc_kernel --launch-skip 505 --launch-count 5 --csv \
    --log-file gpurun_out/r1r_$name.csv python scripts/profile_unet.py > /dev/null 2>&1
}
run base A=1
run nopf RAMP_GEMM_PREFETCH=0
run promo128 RAMP_GEMM_L2PROMO=128
run nocoop RAMP_GEMM_COOP=0
run nocoop_nopf RAMP_GEMM_COOP=0 RAMP_GEMM_PREFETCH=0
# un-profiled timing of the same variants
for v in "A=1" "RAMP_GEMM_PREFETCH=0" "RAMP_GEMM_L2PROMO=128" "RAMP_GEMM_COOP=0"; do
  echo "== $v"; env $v timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['roofline']['class_ms'])"
done > gpurun_out/r1r_variants.txt 2>&1
cat gpurun_out/r1r_variants.txt
