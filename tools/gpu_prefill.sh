#!/bin/bash
# GPU-box script: prefix-prefill kernels — staged probe, parity suite, timing vs the reference's Triton kernels
# (each step under its own timeout: a wedged mbarrier wait must not hang the box).
#   B200PA_PREFILL_TC=0        force the mma.sync family
#   B200PA_PREFILL_WS=0        tcgen05 family without warp specialisation
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 120 python tools/tc_probe.py > gpurun_out/prefill_probe.log 2>&1; echo "exit=$?" >> gpurun_out/prefill_probe.log
if grep -q "exit=0" gpurun_out/prefill_probe.log; then
( timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill.log
fi
cat gpurun_out/prefill_probe.log
tail -n 4 gpurun_out/pytest_prefill.log
cat gpurun_out/bench_prefill.log
