#!/bin/bash
tag=$1
mkdir -p gpurun_out
timeout 200 python profiles/ubench/dbg_resume.py > gpurun_out/${tag}_dbg_resume.txt 2>&1
for l in 0 1 6; do timeout 120 python profiles/ubench/epi_trace16.py $l 12 > gpurun_out/${tag}_trace_$l.txt 2>&1; done
FQLPOP_CHAIN_PAIR=0 timeout 120 python profiles/ubench/epi_trace16.py 0 12 > gpurun_out/${tag}_trace_0_nopair.txt 2>&1
timeout 300 python -m pytest tests/test_update_gpu.py -m gpu -q --timeout 180 2>&1 | tail -4
tail -12 gpurun_out/${tag}_dbg_resume.txt; cat gpurun_out/${tag}_trace_0.txt
