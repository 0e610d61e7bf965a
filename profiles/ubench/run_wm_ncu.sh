#!/bin/bash
tag=$1
mkdir -p gpurun_out
WM_STEPS=2 timeout 300 python profiles/ubench/wm_multistep_time.py > gpurun_out/${tag}_plain.log 2>&1 &&
WM_STEPS=2 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_wm_launches.csv python profiles/ubench/wm_multistep_time.py > gpurun_out/${tag}_ncu.log 2>&1
python profiles/summarize_launches.py gpurun_out/${tag}_wm_launches.csv
