#!/bin/bash
# Round-1 evidence run: GPU test suite, both bench arms, ncu launch list of the bench command, per-launch DRAM traffic
# of every GEMM launch of one network call, and full captures of the dominant launches.
mkdir -p gpurun_out
O=gpurun_out/r1f
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -4 > ${O}_tests.log
timeout 600 python bench.py > ${O}_bench.json 2> ${O}_bench.err
timeout 600 python bench.py --impl reference > ${O}_bench_reference.json 2> ${O}_bench_reference.err
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > ${O}_smoke.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file ${O}_launches.csv \
  python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field > ${O}_ncu_bench.log 2>&1
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  --kernel-name regex:gemm_tc_kernel --launch-skip 260 --launch-count 260 --csv --log-file ${O}_gemm_traffic.csv \
  python scripts/profile_unet.py > ${O}_ncu_traffic.log 2>&1
for spec in e1:268 e2:506 e4ff:507; do
  name=${spec%%:*}; skip=${spec##*:}
  timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:gemm_tc_kernel \
    --launch-skip $skip --launch-count 1 -f -o ${O}_gemm_$name python scripts/profile_unet.py > ${O}_ncu_$name.log 2>&1
done
timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:attn_bwd_tc \
    --launch-skip 20 --launch-count 1 -f -o ${O}_attn_bwd python scripts/profile_unet.py > ${O}_ncu_attn.log 2>&1
cat ${O}_tests.log; cat ${O}_smoke.log | tail -2; head -c 600 ${O}_bench.json; echo; head -c 400 ${O}_bench_reference.json
