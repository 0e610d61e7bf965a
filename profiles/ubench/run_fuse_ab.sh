#!/bin/bash
# fused Adam-in-wgrad-epilogue vs separate adam_kernel: parity tests of the fused mode, then the bench both ways
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_baseline_configs_gpu.py tests/test_update_gpu.py -m gpu -q --timeout 300 2>&1 | tail -3
for fuse in 1 0; do
  FQLPOP_FUSE_ADAM=$fuse timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_fuse$fuse.json 2> gpurun_out/${tag}_fuse$fuse.err
  python - <<P
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_fuse$fuse.json") if l.startswith("{")][-1])
    print("fuse=$fuse", round(d["value"]), "updates/s e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 4), d["timed_blocks_ms"])
except Exception as e: print("fuse=$fuse FAILED", e)
P
done
