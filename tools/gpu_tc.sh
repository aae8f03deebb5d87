#!/bin/bash
# GPU-box script: tcgen05 prefill kernel — probe, parity suite, timing (every step under its own timeout).
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 120 python tools/tc_probe.py > gpurun_out/tc_probe.log 2>&1; echo "exit=$?" >> gpurun_out/tc_probe.log
if grep -q "exit=0" gpurun_out/tc_probe.log; then
( timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_tc.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_tc.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill_tc.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill_tc.log
fi
cat gpurun_out/tc_probe.log
tail -n 4 gpurun_out/pytest_prefill_tc.log
cat gpurun_out/bench_prefill_tc.log
