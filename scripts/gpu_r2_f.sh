#!/bin/bash
# Round-2 run F: the no-allocation entry point's test alone in a fresh process, then run E again (the machine that held run
# E's outputs was replaced): whole GPU suite (the new test runs last), bench with every section, reference arm, smoke, ncu pass.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -3 | tee gpurun_out/r2f_no_alloc_alone.log
bash scripts/gpu_r2_e.sh
