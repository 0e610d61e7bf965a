#!/bin/bash
# rollout parity tests + gen A/B + the rollout16 traces
tag=$1
bash profiles/ubench/run_rollout_gen.sh $tag
FQLPOP_ROLLOUT_PAIR=0 TRACE_STAGES=1 timeout 300 python profiles/ubench/ro_trace16.py 13 26 > gpurun_out/${tag}_ro_trace.txt 2>&1
FQLPOP_ROLLOUT_PAIR=1 TRACE_STAGES=1 timeout 300 python profiles/ubench/ro_trace16.py 13 26 > gpurun_out/${tag}_ro_trace_pair.txt 2>&1
head -28 gpurun_out/${tag}_ro_trace.txt; sed -n 28,60p gpurun_out/${tag}_ro_trace.txt
