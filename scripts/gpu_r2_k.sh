#!/bin/bash
# Round-2 run K: ncu --set full with source import of the GEGLU-forward (e1) and GEGLU-backward (e2) launches at L = 64:
# per-instruction stall samples of the epilogue's store path.
mkdir -p gpurun_out
O=gpurun_out/r2k
NCU="ncu --clock-control none"
export RAMP_UNET_SPLIT=1
python scripts/profile_unet.py > ${O}_unet.plain 2>&1 || { tail -3 ${O}_unet.plain; exit 1; }
full() {  # name kernel-regex skip
  $NCU --set full --import-source on --kernel-name-base demangled -k "regex:$2" -s $3 -c 1 -f -o ${O}_$1 python scripts/profile_unet.py > ${O}_$1.log 2>&1
  echo "$1 rc $?"
  ncu -i ${O}_$1.ncu-rep --page raw --csv > ${O}_$1.raw.csv 2>/dev/null
  ncu -i ${O}_$1.ncu-rep --page source --csv --print-source sass > ${O}_$1.source.csv 2>/dev/null
  rm -f ${O}_$1.ncu-rep
  ls -la ${O}_$1.*
}
full e1 "gemm_tc_kernel<.int.1," 16
full e2 "gemm_tc_kernel<.int.2," 30
