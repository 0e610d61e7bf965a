#!/bin/bash
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_worldmodel_gpu.py -m gpu -q --timeout 300 2>&1 | grep -v "^  \|Warning\|^$" | tail -${TAILN:-30}
FQLPOP_LIB=$PWD/flow-q-learning_b200/lib/libfqlpop_trace.so WM_STEPS=1 timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -3
for scan in 1 0; do
  FQLPOP_WM_SCAN=$scan timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1
done
FQLPOP_WM_SCAN=1 WM_TERM_WEIGHT=0 timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1
