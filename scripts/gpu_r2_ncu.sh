#!/bin/bash
# Round-2 ncu evidence (one gpurun call; every profiled command is first run plain and must exit 0):
#   1. nearest-point kernel, --set full, one launch per dimension at the sweep's shape (8192 x 64 waypoints x 1 M points)
#   2. DRAM bytes of all 260 GEMM launches of one 7168-row score-network call (roofline.traffic)
#   3. --set full of the dominant GEMM launches at L = 64 (GEGLU fwd, FF2^T + GEGLU bwd, FF1^T + LN bwd) and of the
#      attention backward at L = 64
#   4. launch list of the bench command (second eager sampling of the warm-up: every launch with its device time)
mkdir -p gpurun_out
O=gpurun_out/r2n
NCU="ncu --clock-control none"
for d in 2 3; do
  python scripts/profile_field.py $d > ${O}_field$d.plain 2>&1 && \
  $NCU --set full --import-source on -k regex:nn_query_kernel -s 2 -c 1 -f -o ${O}_nn$d python scripts/profile_field.py $d > ${O}_nn$d.log 2>&1
  echo "nn d=$d rc $?"
  ncu -i ${O}_nn$d.ncu-rep --page raw --csv > ${O}_nn$d.raw.csv 2>/dev/null
  ncu -i ${O}_nn$d.ncu-rep --page source --csv > ${O}_nn$d.source.csv 2>/dev/null
  [ $d = 3 ] && rm -f ${O}_nn$d.ncu-rep
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
full attn_fwd64 "attn_fwd_tc<.int.64," 3
unset RAMP_UNET_SPLIT
python bench.py --steps 1 --warmup 3 --cpu-sample 0 --no-sections --eager > ${O}_bench.plain 2> ${O}_bench.plain.err && \
$NCU --metrics gpu__time_duration.sum --launch-skip 7700 --launch-count 7600 --csv --log-file ${O}_launches.csv \
    python bench.py --steps 1 --warmup 3 --cpu-sample 0 --no-sections --eager > ${O}_ncu_bench.log 2>&1
echo "launch list rc $?"
ls -la gpurun_out | grep r2n
