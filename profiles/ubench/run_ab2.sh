#!/bin/bash
# GPU-box script: all GPU tests, then the default bench (update step only) and the chain roofline launches.
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 180 > gpurun_out/${tag}_pytest_all.log 2>&1
tail -5 gpurun_out/${tag}_pytest_all.log
timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
FQLPOP_CHAIN_PAIR=0 timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-worldmodel --no-rollout > gpurun_out/${tag}_bench_nopair.json 2> gpurun_out/${tag}_bench_nopair.err
python - <<P
import json
for k in ("bench", "bench_nopair"):
    try:
        d = json.loads([l for l in open("gpurun_out/${tag}_%s.json" % k) if l.startswith("{")][-1])
        print(k, round(d["value"]), "updates/s e2e", round(d["e2e"]["value"]), "ms/step", d["ms_per_step"], "roofline", d["roofline"]["frac"], d["roofline"].get("launch_us"))
        print("    policy pass launch_ms", d["roofline"]["launch_ms"])
        for k, r in d.get("roofline_other_launches", {}).items(): print("   ", k, round(r["launch_ms"], 4), round(r["frac"], 3))
    except Exception as e: print(k, "FAILED", e)
P
tail -3 gpurun_out/${tag}_bench.err
