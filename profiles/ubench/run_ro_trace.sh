#!/bin/bash
tag=$1
mkdir -p gpurun_out
timeout 120 ./profiles/ubench/mma_rate/mma_rate | tail -4
TRACE_STAGES=1 timeout 300 python profiles/ubench/ro_trace16.py 13 26 > gpurun_out/${tag}_ro_trace.txt 2>&1
FQLPOP_ROLLOUT_PAIR=1 TRACE_STAGES=1 timeout 300 python profiles/ubench/ro_trace16.py 13 26 > gpurun_out/${tag}_ro_trace_pair.txt 2>&1
head -16 gpurun_out/${tag}_ro_trace.txt
sed -n 28,95p gpurun_out/${tag}_ro_trace_pair.txt
