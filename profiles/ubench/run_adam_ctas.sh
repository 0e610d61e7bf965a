#!/bin/bash
tag=$1
mkdir -p gpurun_out
for c in 16 24 32 48 64; do
  FQLPOP_ADAM_CTAS=$c timeout 300 python bench.py --steps 200 --warmup 20 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_c$c.json 2> gpurun_out/${tag}_c$c.err
  python - <<P
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_c$c.json") if l.startswith("{")][-1])
    print("adam ctas $c:", round(d["value"]), "updates/s, ms/step", round(d["ms_per_step"], 4))
except Exception as e: print("ctas=$c FAILED", e)
P
done
