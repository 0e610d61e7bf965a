#!/bin/bash
# GPU-box A/B script: parity tests, bench with the second-generation chain kernels vs the first, per-CTA timeline.
# usage: bash profiles/ubench/run_ab.sh <tag>
tag=${1:-r02a}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_update_gpu.py tests/test_gemm_gpu.py -m gpu -x -q --timeout 180 > gpurun_out/${tag}_pytest_update.log 2>&1
echo "rc=$?" >> gpurun_out/${tag}_pytest_update.log
tail -5 gpurun_out/${tag}_pytest_update.log
timeout 300 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/${tag}_pytest_all.log 2>&1
tail -5 gpurun_out/${tag}_pytest_all.log
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
FQLPOP_CHAIN_GEN=1 timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_bench_gen1.json 2> gpurun_out/${tag}_bench_gen1.err
timeout 120 python profiles/ubench/timeline.py 16 > gpurun_out/${tag}_timeline.txt 2>&1
tail -3 gpurun_out/${tag}_timeline.txt
head -c 700 gpurun_out/${tag}_bench.json; echo
head -c 300 gpurun_out/${tag}_bench_gen1.json; echo
