#!/bin/bash
tag=$1
mkdir -p gpurun_out
B="python bench.py --steps 100 --warmup 10 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout"
run() { name=$1; shift; env "$@" timeout 200 $B > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; python - <<P
import json
try:
    d=json.loads([l for l in open("gpurun_out/${tag}_$name.json") if l.startswith("{")][-1]); print("$name", round(d["value"]), round(d["e2e"]["value"]), d["ms_per_step"], d["gpu_launches"])
except Exception as e: print("$name", "FAILED", e)
P
}
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/${tag}_pytest_all.log 2>&1; tail -8 gpurun_out/${tag}_pytest_all.log
run fused X=1
run unfused FQLPOP_FUSE_ADAM=0
run fused_g2 FQLPOP_GROUP_SIZE=2
run fused_g8 FQLPOP_GROUP_SIZE=8
