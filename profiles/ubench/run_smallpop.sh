#!/bin/bash
# small populations (late successive-halving rungs): second-generation pair kernel vs first-generation N-split cluster kernel
tag=$1
mkdir -p gpurun_out
for P in 1 2 4 8; do
for mode in gen2 chain2 gen2nopair; do
  case $mode in gen2) ev="X=1";; chain2) ev="FQLPOP_CHAIN2=1";; gen2nopair) ev="FQLPOP_CHAIN_PAIR=0";; esac
  env $ev timeout 200 python bench.py --population $P --steps 200 --warmup 20 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_p${P}_$mode.json 2> gpurun_out/${tag}_p${P}_$mode.err
  python - <<P2
import json
try:
    d=json.loads([l for l in open("gpurun_out/${tag}_p${P}_$mode.json") if l.startswith("{")][-1]); print("P=$P $mode", round(d["value"]), d["ms_per_step"])
except Exception as e: print("P=$P $mode FAILED", e)
P2
done; done
