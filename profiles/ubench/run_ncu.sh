#!/bin/bash
# usage: bash profiles/ubench/run_ncu.sh <tag> <which> [kernel-regex]: ncu --set full of one debug chain launch (3 launches; the last 2 captured)
tag=$1; which=$2; rx=${3:-fql_chain}
mkdir -p gpurun_out
python profiles/ubench/roofline_launch.py $which > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 121 --launch-count 2 -o gpurun_out/${tag}_w${which} -f python profiles/ubench/roofline_launch.py $which > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_ncu.log
