#!/bin/bash
# usage: bash profiles/ubench/run_trace.sh <tag> : traces of the main chain launches + no-weight-stream A/B of the Euler chain
tag=${1:-r02c}
mkdir -p gpurun_out
for l in 0 1 3 4 6; do timeout 120 python profiles/ubench/epi_trace16.py $l 12 > gpurun_out/${tag}_trace_$l.txt 2>&1; done
FQLPOP_LIB=$PWD/flow-q-learning_b200/lib/libfqlpop_trace.so timeout 120 python profiles/ubench/weight_stream.py > gpurun_out/${tag}_ws.txt 2>&1
FQLPOP_CHAIN_DEBUG=1 FQLPOP_LIB=$PWD/flow-q-learning_b200/lib/libfqlpop_trace.so timeout 120 python profiles/ubench/weight_stream.py >> gpurun_out/${tag}_ws.txt 2>&1
FQLPOP_CHAIN_DEBUG=1 timeout 120 python profiles/ubench/epi_trace16.py 0 12 > gpurun_out/${tag}_trace_0_nows.txt 2>&1
cat gpurun_out/${tag}_trace_0.txt gpurun_out/${tag}_ws.txt
