#!/bin/bash
tag=$1
mkdir -p gpurun_out
B="python bench.py --steps 100 --warmup 10 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout"
run() { name=$1; shift; env "$@" timeout 200 $B > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; python - <<P
import json
try:
    d=json.loads([l for l in open("gpurun_out/${tag}_$name.json") if l.startswith("{")][-1]); print("$name", round(d["value"]), round(d["e2e"]["value"]), d["ms_per_step"])
except Exception as e: print("$name", "FAILED", e)
P
}
run default X=1
run euler_chain2 FQLPOP_CHAIN2=2
run euler_chain2_g2 FQLPOP_CHAIN2=2 FQLPOP_GROUP_SIZE=2
run euler_chain2_g8 FQLPOP_CHAIN2=2 FQLPOP_GROUP_SIZE=8
run all_chain2 FQLPOP_CHAIN2=1
B="$B --population 12"; run p12_default X=1; run p12_euler_chain2 FQLPOP_CHAIN2=2; run p12_chain2 FQLPOP_CHAIN2=1
