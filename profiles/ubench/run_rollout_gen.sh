#!/bin/bash
# A/B of the two whole-rollout kernels (FQLPOP_ROLLOUT_GEN) x pair mode, after the rollout parity tests.
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_rollout_gpu.py tests/test_host_gpu.py "tests/test_baseline_configs_gpu.py::test_config3_rollout_at_scale_matches_oracle" -m gpu -q --timeout 300 2>&1 | tail -12
for mode in gen2 gen2nopair gen1; do
  unset FQLPOP_ROLLOUT_PAIR FQLPOP_ROLLOUT_GEN
  if [ $mode = gen2nopair ]; then export FQLPOP_ROLLOUT_PAIR=0; fi
  if [ $mode = gen1 ]; then export FQLPOP_ROLLOUT_GEN=1; fi
  timeout 300 python bench.py --steps 20 --warmup 5 --blocks 1 --no-cpu-baseline --no-worldmodel > gpurun_out/${tag}_$mode.json 2> gpurun_out/${tag}_$mode.err
  python - <<P
import json
try:
    d=json.loads([l for l in open("gpurun_out/${tag}_$mode.json") if l.startswith("{")][-1]); r=d["rollout"]
    print("$mode", round(r["value"]/1e6,1), "M env-steps/s", r["ms_per_evaluation"], "ms; no-termination:", round(r["no_termination"]["value"]/1e6,1), r["mean_success_rate"], r["lockstep_steps_per_member"])
except Exception as e: print("$mode FAILED", e)
P
done
