#!/bin/bash
# 2 GPUs: device round trip of a member, the 2-rank Trainer test with migration over NCCL, then config 4 with and without
cd /root/repo
timeout 600 python -m pytest tests/test_migration_gpu.py tests/test_dist_gpu.py -x -q 2>&1 | tail -15 > gpurun_out/migrate_tests.log
for rb in 1 0; do
  FQLPOP_REBALANCE=$rb timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
    --master-port 29611 bench.py --gpus 2 --workload sh64 > gpurun_out/sh64_2gpu_rb$rb.json 2> gpurun_out/sh64_2gpu_rb$rb.err
done
