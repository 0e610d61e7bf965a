#!/bin/bash
# ncu --set full captures for profiles/: the roofline launch (policy forward pass), the critic LayerNorm forward and backward
# launches, the Euler chain, adam_kernel and prep_kernel inside a graph-replayed step. Each only after the same command
# exited 0 without ncu. usage: run_ncu_all.sh <tag>
tag=$1
mkdir -p gpurun_out
for w in 1 2 3; do
python profiles/ubench/roofline_launch.py $w > gpurun_out/${tag}_plain_$w.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fql_chain --launch-skip 121 --launch-count 2 -o gpurun_out/${tag}_chain_w$w -f python profiles/ubench/roofline_launch.py $w > gpurun_out/${tag}_ncu_$w.log 2>&1
tail -1 gpurun_out/${tag}_ncu_$w.log
done
python profiles/ubench/roofline_launch.py 1 > gpurun_out/${tag}_plain_step.log 2>&1 &&
ncu --set full --clock-control none -k regex:'adam_kernel|prep_kernel|fql_gemm_kernel' --launch-skip 24 --launch-count 6 -o gpurun_out/${tag}_step_kernels -f python profiles/ubench/roofline_launch.py 1 > gpurun_out/${tag}_ncu_step.log 2>&1
tail -1 gpurun_out/${tag}_ncu_step.log
# launch list of the bench command (shares of the step)
python bench.py --steps 5 --warmup 3 --blocks 1 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 5 --warmup 3 --blocks 1 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_ncu_bench.log 2>&1
tail -2 gpurun_out/${tag}_ncu_bench.log | cut -c1-200
