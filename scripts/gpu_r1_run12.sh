set -x
for v in v0 v1 v2; do
RAMP_B200_LIB=$PWD/ramp_b200/variants/lib_$v.so timeout 300 python bench.py --steps 1 --warmup 3 --cpu-sample 0 --replan-ticks 0 --no-field --max-rows 7168 --dump-launches gpurun_out/r1l_events_$v.csv > gpurun_out/r1l_bench_$v.json 2> gpurun_out/r1l_bench_$v.err
done
timeout 200 python - > gpurun_out/r1l_field.json 2> gpurun_out/r1l_field.err <<'PY'
import json, sys, torch
sys.path.insert(0, '.')
import bench
print(json.dumps(bench.field_roofline(torch.device('cuda:0'), 1965)))
PY
