#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_sampler_gpu.py tests/test_fullsize_gpu.py -q -k "nn_query" 2>&1 | tail -2
{
for dim in 2 3; do
  echo "auto:"; timeout 120 python scripts/time_field.py $dim
  for t in 1024 2048; do for r in 4 6 8 10 12; do
    echo -n "R=$r TILE=$t: "; RAMP_NN_FORCE=$r,$t timeout 120 python scripts/time_field.py $dim
  done; done
done
echo "small batch (replan shape):"; timeout 120 python scripts/time_field.py 2 1024
} > gpurun_out/nn_sweep.txt 2>&1
cat gpurun_out/nn_sweep.txt
