#!/bin/bash
# Round-3 evidence: default bench line, ncu --set full of the rollout16 kernel and of the roofline launch (policy forward
# pass), launch list of the bench command. Every ncu run only after the same command exited 0 without ncu.
tag=$1
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python profiles/ubench/rollout_once.py > gpurun_out/${tag}_plain_ro.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fql_rollout16 --launch-skip 1 --launch-count 1 -o gpurun_out/${tag}_rollout16 -f python profiles/ubench/rollout_once.py > gpurun_out/${tag}_ncu_ro.log 2>&1
tail -2 gpurun_out/${tag}_ncu_ro.log
python profiles/ubench/roofline_launch.py 1 > gpurun_out/${tag}_plain_1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fql_chain --launch-skip 121 --launch-count 2 -o gpurun_out/${tag}_chain_w1 -f python profiles/ubench/roofline_launch.py 1 > gpurun_out/${tag}_ncu_1.log 2>&1
tail -1 gpurun_out/${tag}_ncu_1.log
python bench.py --steps 5 --warmup 3 --blocks 1 --no-cpu-baseline --no-worldmodel > gpurun_out/${tag}_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 5 --warmup 3 --blocks 1 --no-cpu-baseline --no-worldmodel > gpurun_out/${tag}_ncu_bench.log 2>&1
tail -2 gpurun_out/${tag}_ncu_bench.log | cut -c1-200
head -c 1500 gpurun_out/${tag}_bench.json
