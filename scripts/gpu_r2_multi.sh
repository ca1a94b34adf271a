#!/bin/bash
# Round-2 multi-GPU evidence.  usage: gpurun --gpus G -- 'bash scripts/gpu_r2_multi.sh "<N list>" [headline]'
#   sweep (BASELINE config 5, STRONG scaling): 65,536 trajectories x 1 M-point cloud, 65,536 / N per GPU, for each N in the list
#   headline (weak scaling, 1024 per GPU) at the largest N, with per-rank times and the all-gather timed
mkdir -p gpurun_out
O=gpurun_out/r2m
PORT=29511
for N in $1; do
  if [ $N = 1 ]; then
    timeout 900 python bench.py --workload sweep --gpus 1 --steps 1 --warmup 1 > ${O}_sweep_n1.json 2> ${O}_sweep_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --workload sweep --gpus $N --steps ${STEPS:-2} --warmup 1 > ${O}_sweep_n$N.json 2> ${O}_sweep_n$N.err
  fi
  echo "sweep N=$N rc $?"; tail -c 1200 ${O}_sweep_n$N.json | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N', d['n_gpus'], 'traj/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],1), 'per-rank', [round(x) for x in d['per_rank_ms_per_step']], 'nn share', round(d['roofline']['share_of_step'],3), 'nn frac', round(d['roofline']['frac'],3), d['clocks'])
except Exception as e: print('parse failed', e)
"
  PORT=$((PORT+1))
done
if [ -n "$2" ]; then
  N=$(echo $1 | awk '{print $NF}')
  for mode in "" "--eager"; do
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --gpus $N --steps 5 --warmup 3 --time-gather $mode > ${O}_headline_n$N$mode.json 2> ${O}_headline_n$N$mode.err
    echo "headline N=$N $mode rc $?"; python -c "
import json
try:
    d=json.loads(open('${O}_headline_n$N$mode.json').read().strip().splitlines()[-1]); print('N', d['n_gpus'], 'traj/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],1), 'per-rank', [round(x) for x in d['per_rank_ms_per_step']], 'allgather ms', d['allgather_ms'], d['clocks'])
except Exception as e: print('parse failed', e)
"
    PORT=$((PORT+1))
  done
fi
