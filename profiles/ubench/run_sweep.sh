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
export FQLPOP_ADAM_STREAM=0 FQLPOP_STAGGER_US=0 FQLPOP_LIB=$PWD/flow-q-learning_b200/lib/libfqlpop_trace.so
run tracelib X=1
run skip_adam FQLPOP_DEBUG_SKIP=1
run skip_adam_wgrad FQLPOP_DEBUG_SKIP=3
run skip_all7 FQLPOP_DEBUG_SKIP=7
run noweights FQLPOP_CHAIN_DEBUG=1
run noweights_skip7 FQLPOP_CHAIN_DEBUG=1 FQLPOP_DEBUG_SKIP=7
run skip7_g2 FQLPOP_DEBUG_SKIP=7 FQLPOP_GROUP_SIZE=2
B="$B --population 32"; run p32 X=1
