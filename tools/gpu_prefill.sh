#!/bin/bash
# GPU-box script: prefix-prefill parity suite + timing vs the reference's Triton kernels.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill.log
( B200PA_PREFILL_MT=1 timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q -k "realistic or oracle" ) > gpurun_out/pytest_prefill_mt1.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_mt1.log
timeout 900 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill.log
B200PA_PREFILL_MT=1 timeout 900 python tools/bench_prefill.py > gpurun_out/bench_prefill_mt1.log 2>&1
timeout 300 python tools/bench_cache_ops.py > gpurun_out/cache_ops.log 2>&1
tail -n 6 gpurun_out/pytest_prefill.log gpurun_out/pytest_prefill_mt1.log
cat gpurun_out/bench_prefill.log gpurun_out/bench_prefill_mt1.log
grep convert_fp8 gpurun_out/cache_ops.log
