#!/bin/bash
# usage: run_bench_tl.sh <tag>: update tests, bench, 5-step occupancy profile
tag=$1
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_update_gpu.py tests/test_host_gpu.py -m gpu -q --timeout 180 2>&1 | tail -5 > gpurun_out/${tag}_pytest.log; cat gpurun_out/${tag}_pytest.log
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
head -c 300 gpurun_out/${tag}_bench.json; echo
TL_STEPS=5 TL_PROFILE=50 timeout 120 python profiles/ubench/timeline.py 16 > gpurun_out/${tag}_timeline.txt 2>&1
grep -A95 "occupancy profile" gpurun_out/${tag}_timeline.txt | awk 'NR<3 || NR%2==0' | head -60
