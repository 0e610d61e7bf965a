#!/bin/bash
# end-of-round check: the whole GPU suite, smoke(), then the default bench line
cd /root/repo
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > gpurun_out/final_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
