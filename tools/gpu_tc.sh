#!/bin/bash
# GPU-box script: tcgen05 prefill kernel — parity suite with the family forced on / off, then timing.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 120 python tools/tc_probe.py > gpurun_out/tc_probe.log 2>&1; echo "exit=$?" >> gpurun_out/tc_probe.log
( timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_tc.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_tc.log
( B200PA_PREFILL_TC=0 timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_mma.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_mma.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill_tc.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill_tc.log
tail -n 4 gpurun_out/tc_probe.log gpurun_out/pytest_prefill_tc.log gpurun_out/pytest_prefill_mma.log
cat gpurun_out/bench_prefill_tc.log
