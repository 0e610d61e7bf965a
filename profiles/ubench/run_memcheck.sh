#!/bin/bash
# compute-sanitizer memcheck of smoke() (update step, rollout) and of the small rollout / no-alloc / world-model scan tests
tag=$1
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 --log-file gpurun_out/${tag}_memcheck_smoke.log python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${tag}_smoke.out 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/${tag}_smoke.out; tail -4 gpurun_out/${tag}_memcheck_smoke.log
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 --log-file gpurun_out/${tag}_memcheck_tests.log python -m pytest tests/test_zz_no_alloc.py "tests/test_rollout_gpu.py::test_device_philox_stream_is_bit_exact_and_gather_uses_it" "tests/test_worldmodel_gpu.py::test_multistep_predictor_steps_match_oracle" -m gpu -x -q --timeout 1200 > gpurun_out/${tag}_tests.out 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/${tag}_tests.out; tail -4 gpurun_out/${tag}_memcheck_tests.log
