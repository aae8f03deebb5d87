#!/bin/bash
# GPU-box script: prefix-prefill parity suite + timing vs the reference's Triton kernels.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill.log
timeout 900 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill.log
tail -n 15 gpurun_out/pytest_prefill.log
cat gpurun_out/bench_prefill.log
