#!/bin/bash
# 8-GPU box, end of round 2: the 2-GPU NCCL test, the driver-style default line at N=8 (rollout + sh64 + CPU baseline ride
# along), BASELINE config 5 at N=8. usage: run_8gpu_r03.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout 500 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 900 $TR bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/${tag}_default_8gpu.json 2> gpurun_out/${tag}_default_8gpu.err
echo "default rc=$?"; tail -2 gpurun_out/${tag}_default_8gpu.err | cut -c1-300
timeout 600 $TR bench.py --gpus 8 --population 32 --batch 1024 --shape cube-double --steps 20 --warmup 5 --blocks 3 --no-cpu-baseline --no-worldmodel --no-sh64 --no-rollout > gpurun_out/${tag}_config5_8gpu.json 2> gpurun_out/${tag}_config5_8gpu.err
echo "config5 rc=$?"; tail -2 gpurun_out/${tag}_config5_8gpu.err | cut -c1-300
python - <<P
import json
for f in ("default","config5"):
    try:
        d=json.loads([l for l in open("gpurun_out/${tag}_%s_8gpu.json"%f) if l.startswith("{")][-1])
        print(f, round(d["value"]), "e2e", round(d["e2e"]["value"]), d["ms_per_step"], d.get("hbm_bytes_in_use_rank0"))
        r=d.get("rollout")
        if r: print(" rollout", round(r["value"]/1e6,1), "M env-steps/s, e2e", round(r["e2e"]["value"]/1e6,1))
        sh=d.get("sh64")
        if sh:
            print(" sh64", round(sh["value"]), sh["total_s"])
            for r in sh["per_rung"]: print("  ", {k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items()})
    except Exception as e: print(f, "FAILED", e)
P
