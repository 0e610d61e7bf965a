#!/bin/bash
# 8-GPU box: BASELINE config 5 (256 alpha = 32 per GPU, batch 1024, cube-double) with config 4 (sh64) riding along,
# then the default config-2 line at N=8 for reference. usage: run_8gpu.sh <tag>
tag=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 600 $TR bench.py --gpus 8 --population 32 --batch 1024 --shape cube-double --steps 20 --warmup 5 --blocks 3 --no-cpu-baseline --no-worldmodel > gpurun_out/${tag}_config5_8gpu.json 2> gpurun_out/${tag}_config5_8gpu.err
tail -2 gpurun_out/${tag}_config5_8gpu.err
timeout 600 $TR bench.py --gpus 8 --steps 100 --warmup 10 --blocks 3 --no-cpu-baseline --no-worldmodel --no-sh64 --no-rollout > gpurun_out/${tag}_config2_8gpu.json 2> gpurun_out/${tag}_config2_8gpu.err
tail -2 gpurun_out/${tag}_config2_8gpu.err
python - <<P
import json
for f in ("config5","config2"):
    try:
        d=json.loads([l for l in open("gpurun_out/${tag}_%s_8gpu.json"%f) if l.startswith("{")][-1])
        print(f, round(d["value"]), d["ms_per_step"], d.get("hbm_bytes_in_use_rank0"))
        sh=d.get("sh64")
        if sh:
            print(" sh64", round(sh["value"]), sh["total_s"])
            for r in sh["per_rung"]: print("  ", {k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items()})
    except Exception as e: print(f, "FAILED", e)
P
