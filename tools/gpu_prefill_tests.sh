#!/bin/bash
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_prefill.py -m gpu -q -k "random or deterministic or long_context" ) > gpurun_out/pytest_prefill_new.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_new.log
( B200PA_PREFILL_TC=0 timeout 1200 python -m pytest tests/test_gpu_prefill.py -m gpu -q -k "random or deterministic or long_context" ) > gpurun_out/pytest_prefill_new_mma.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_new_mma.log
tail -n 30 gpurun_out/pytest_prefill_new.log; tail -n 8 gpurun_out/pytest_prefill_new_mma.log
