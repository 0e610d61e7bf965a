#!/bin/bash
tag=$1
mkdir -p gpurun_out
WM_STEPS=1 timeout 300 python profiles/ubench/wm_multistep_time.py > gpurun_out/${tag}_plain.log 2>&1 &&
WM_STEPS=1 ncu --set full --clock-control none --import-source on -k regex:wm_scan --launch-skip 6 --launch-count 2 -o gpurun_out/${tag}_wm_scan -f python profiles/ubench/wm_multistep_time.py > gpurun_out/${tag}_ncu.log 2>&1
tail -2 gpurun_out/${tag}_ncu.log
