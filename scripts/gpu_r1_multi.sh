#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field > gpurun_out/r1_multi_n$N.json 2> gpurun_out/r1_multi_n$N.err
tail -c 1500 gpurun_out/r1_multi_n$N.json; tail -5 gpurun_out/r1_multi_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/r1_multi_ref_n$N.json 2> gpurun_out/r1_multi_ref_n$N.err
tail -c 600 gpurun_out/r1_multi_ref_n$N.json
