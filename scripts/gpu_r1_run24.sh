#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r1y
timeout 300 python -m pytest tests/test_unet_gpu.py -x -q 2>&1 | tail -3
timeout 400 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --dump-launches ${O}_events.csv > ${O}_bench.json 2> ${O}_bench.err
python -c "
import json
d=json.load(open('${O}_bench.json'))
print(d['value'],d['ms_per_step'],d['roofline']['class_ms'])"
timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:attn_bwd_tc --launch-skip 30 --launch-count 1 -f -o ${O}_attnb64 python scripts/profile_unet.py > ${O}_ncu_attnb64.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:gn_backward --launch-skip 47 --launch-count 1 -f -o ${O}_gnb_c32_v2 python scripts/profile_unet.py > ${O}_ncu_gnb2.log 2>&1
