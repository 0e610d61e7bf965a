#!/bin/bash
# multistep world-model trainer: parity tests, then the step time by number of time chunks
cd /root/repo
timeout 600 python -m pytest tests/test_worldmodel_gpu.py -m gpu -q -x --timeout 300 2>&1 | tail -2
for c in ${CHUNKS:-8 4 1}; do
  FQLPOP_WM_CHUNKS=$c timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1 | sed "s/^/chunks=$c /"
  FQLPOP_WM_CHUNKS=$c WM_TERM_WEIGHT=0 timeout 300 python profiles/ubench/wm_multistep_time.py 2>&1 | tail -1 | sed "s/^/chunks=$c /"
done
