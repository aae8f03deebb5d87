#!/bin/bash
# GPU-box script: phase trace of the warp-specialised prefill kernel (library built with -DB200PA_PREFILL_TRACE),
# the prefill parity file and the prefill timings on the same build.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 300 python tools/prefill_trace.py > gpurun_out/prefill_trace.log 2>&1; echo "exit=$?" >> gpurun_out/prefill_trace.log
cat gpurun_out/prefill_trace.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill_tracebuild.log 2>&1
cat gpurun_out/bench_prefill_tracebuild.log
( time timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill.log
tail -n 12 gpurun_out/pytest_prefill.log
