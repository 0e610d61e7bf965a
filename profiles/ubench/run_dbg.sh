#!/bin/bash
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_update_gpu.py tests/test_baseline_configs_gpu.py -m gpu -q --timeout 300 2>&1 | tail -4
for l in 0 2 3 4 7; do timeout 120 python profiles/ubench/epi_trace16.py $l 5 > gpurun_out/${tag}_trace_$l.txt 2>&1; tail -7 gpurun_out/${tag}_trace_$l.txt; done
timeout 200 python bench.py --steps 100 --warmup 10 --blocks 3 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
python - <<P
import json
d=json.loads([l for l in open("gpurun_out/${tag}_bench.json") if l.startswith("{")][-1]); print("bench", round(d["value"]), round(d["e2e"]["value"]), d["ms_per_step"], d["roofline"]["frac"], {k:round(v["frac"],3) for k,v in d["roofline_other_launches"].items()})
P
TL_STEPS=1 timeout 120 python profiles/ubench/timeline.py 16 > gpurun_out/${tag}_timeline.txt 2>&1; head -12 gpurun_out/${tag}_timeline.txt; tail -1 gpurun_out/${tag}_timeline.txt
