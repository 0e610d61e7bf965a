#!/bin/bash
# world-model trainer: parity tests, then the multistep step time with / without the 64 x 64 GEMM and the scan kernels
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_worldmodel_gpu.py -m gpu -q --timeout 300 2>&1 | grep -E "^E  |FAILED|passed|failed" | head -${TAILN:-8}
for g in 1 0; do
  FQLPOP_WM_GEMM64=$g timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1 | sed "s/^/gemm64=$g /"
  FQLPOP_WM_GEMM64=$g WM_TERM_WEIGHT=0 timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1 | sed "s/^/gemm64=$g /"
done
FQLPOP_WM_SCAN=0 timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1
