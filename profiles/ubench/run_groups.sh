#!/bin/bash
tag=$1
mkdir -p gpurun_out
for gs in 2 4 8 16; do
  FQLPOP_GROUP_SIZE=$gs timeout 300 python bench.py --steps 200 --warmup 20 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_g$gs.json 2> gpurun_out/${tag}_g$gs.err
  python - <<P
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_g$gs.json") if l.startswith("{")][-1])
    print("group size $gs:", round(d["value"]), "updates/s, e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 4))
except Exception as e: print("gs=$gs FAILED", e)
P
done
for P in 24 32; do
  timeout 300 python bench.py --population $P --steps 100 --warmup 10 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_p$P.json 2> gpurun_out/${tag}_p$P.err
  python - <<Q
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_p$P.json") if l.startswith("{")][-1])
    print("population $P:", round(d["value"]), "updates/s, ms/step", round(d["ms_per_step"], 4))
except Exception as e: print("P=$P FAILED", e)
Q
done
