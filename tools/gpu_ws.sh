#!/bin/bash
# GPU-box script: bring-up of the warp-specialised tcgen05 prefill kernel (each step under its own timeout).
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export B200PA_PREFILL_WS=${B200PA_PREFILL_WS:-1}
timeout 120 python tools/tc_probe.py > gpurun_out/ws_probe.log 2>&1; echo "exit=$?" >> gpurun_out/ws_probe.log
if grep -q "exit=0" gpurun_out/ws_probe.log; then
( timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_ws.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_ws.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill_ws.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill_ws.log
fi
cat gpurun_out/ws_probe.log
tail -n 4 gpurun_out/pytest_prefill_ws.log
cat gpurun_out/bench_prefill_ws.log
