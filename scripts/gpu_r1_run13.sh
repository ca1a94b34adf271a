set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1m_tests.log
timeout 400 python bench.py --steps 2 --warmup 3 --cpu-sample 0 --replan-ticks 30 --dump-launches gpurun_out/r1m_events.csv > gpurun_out/r1m_bench.json 2> gpurun_out/r1m_bench.err
