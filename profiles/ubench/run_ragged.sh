#!/bin/bash
# chain16 with ragged / odd row-tile counts and config 5's batch, against the oracle
cd /root/repo
timeout 900 python -m pytest tests/test_baseline_configs_gpu.py -x -q -k "default_kernels" 2>&1 | tail -15 > gpurun_out/ragged.log
