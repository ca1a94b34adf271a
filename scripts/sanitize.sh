#!/bin/bash
# compute-sanitizer pass over the hand-rolled mbarrier / TMEM / smem-staging kernels.
# ONE tool per gpurun call (B200_PROFILING.md: several sanitizer tools in one call have left GPUs unusable):
#     gpurun -- 'bash scripts/sanitize.sh memcheck'      out-of-bounds / misaligned global + shared accesses
#     gpurun -- 'bash scripts/sanitize.sh racecheck'     shared-memory hazards (epilogue staging tiles, named barriers)
# over the GEMM instantiations (tests/test_gemm_gpu.py), one score-network evaluation (tests/test_unet_gpu.py) and the
# nearest-point kernel (tests/test_sampler_gpu.py::test_nn_query_exact).
# Output: gpurun_out/sanitize_<tool>_*.log (summaries are copied under profiles/).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # name timeout pytest-args...
  local name=$1 to=$2; shift 2
  timeout $to $CS --tool $TOOL --error-exitcode 1 --print-limit 20 --launch-timeout 120 \
      python -m pytest "$@" -m gpu -q -x -p no:cacheprovider > gpurun_out/sanitize_${TOOL}_${name}.log 2>&1
  echo "$name $TOOL rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' gpurun_out/sanitize_${TOOL}_${name}.log | tr '\n' ' ')"
}
# plain runs first: the sanitizer only sees programs that exit 0 without it
python -m pytest tests/test_gemm_gpu.py "tests/test_unet_gpu.py::test_energy_grad_matches_oracle" "tests/test_sampler_gpu.py::test_nn_query_exact" -m gpu -q -x > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; exit 1; }
run gemm 1200 tests/test_gemm_gpu.py
run nn 600 "tests/test_sampler_gpu.py::test_nn_query_exact"
run unet 1500 "tests/test_unet_gpu.py::test_energy_grad_matches_oracle"
