#!/bin/bash
# 2-GPU box: all GPU tests (incl. the torchrun NCCL test), bench at N=2 (sh64 rides along), sh64 alone at N=1 for reference
tag=${1:-r02e}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/${tag}_pytest_all.log 2>&1
tail -6 gpurun_out/${tag}_pytest_all.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline --no-worldmodel > gpurun_out/${tag}_bench_2gpu.json 2> gpurun_out/${tag}_bench_2gpu.err
tail -3 gpurun_out/${tag}_bench_2gpu.err
timeout 600 python bench.py --workload sh64 > gpurun_out/${tag}_sh64_1gpu.json 2> gpurun_out/${tag}_sh64_1gpu.err
tail -3 gpurun_out/${tag}_sh64_1gpu.err
head -c 400 gpurun_out/${tag}_bench_2gpu.json; echo; head -c 300 gpurun_out/${tag}_sh64_1gpu.json
