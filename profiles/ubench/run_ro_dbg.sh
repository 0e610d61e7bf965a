#!/bin/bash
# which resource bounds rollout16? (trace build) 0 = normal, 1 = no MMAs, 2 = no weight TMA; pair and non-pair
tag=$1
mkdir -p gpurun_out
for pair in 0 1; do for dbg in 2 1; do
  echo "== pair $pair debug $dbg"
  TRACE_STAGES=1 FQLPOP_ROLLOUT_PAIR=$pair FQLPOP_RO_DEBUG=$dbg timeout 300 python profiles/ubench/ro_trace16.py 26 4 > gpurun_out/${tag}_dbg_${pair}_${dbg}.txt 2>&1
  sed -n 2,4p gpurun_out/${tag}_dbg_${pair}_${dbg}.txt | cut -c1-80
  sed -n 8,44p gpurun_out/${tag}_dbg_${pair}_${dbg}.txt | cut -c1-60
done; done
