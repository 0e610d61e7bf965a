#!/bin/bash
tag=$1
mkdir -p gpurun_out
for l in 0 1 3 4; do timeout 120 python profiles/ubench/epi_trace16.py $l 12 > gpurun_out/${tag}_trace_$l.txt 2>&1; done
for l in 0 1 3 4; do echo "== launch $l"; tail -16 gpurun_out/${tag}_trace_$l.txt; done
