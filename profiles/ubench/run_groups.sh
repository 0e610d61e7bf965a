#!/bin/bash
# FQLPOP_GROUP_SIZE sweep (one split for every call): value = K-step calls, e2e = single-step calls with host batches
tag=$1
mkdir -p gpurun_out
for gs in ${SIZES:-1 2 3 4 8 16}; do
  FQLPOP_GROUP_SIZE=$gs timeout 300 python bench.py --steps 200 --warmup 20 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_g$gs.json 2> gpurun_out/${tag}_g$gs.err
  python - <<P
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_g$gs.json") if l.startswith("{")][-1])
    print("group size $gs:", round(d["value"]), "updates/s, e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 4))
except Exception as e: print("gs=$gs FAILED", e)
P
done
