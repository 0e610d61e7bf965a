#!/bin/bash
# config 4 on 8 GPUs with and without survivor migration
cd /root/repo
for rb in 1 0; do
  FQLPOP_REBALANCE=$rb timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
    --master-port 29621 bench.py --gpus 8 --workload sh64 > gpurun_out/sh64_8gpu_rb$rb.json 2> gpurun_out/sh64_8gpu_rb$rb.err
done
