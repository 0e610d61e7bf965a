#!/bin/bash
# small populations with the default kernel choice (N-split cluster kernel up to 8 members): GPU tests, then updates/s
tag=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -2
for P in 1 2 4 8 12; do
  timeout 200 python bench.py --population $P --steps 200 --warmup 20 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_p${P}.json 2> gpurun_out/${tag}_p${P}.err
  python - <<P2
import json
try:
    d=json.loads([l for l in open("gpurun_out/${tag}_p${P}.json") if l.startswith("{")][-1]); print("P=$P", round(d["value"]), "updates/s, e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 4))
except Exception as e: print("P=$P FAILED", e)
P2
done
