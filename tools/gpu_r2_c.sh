#!/bin/bash
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_gpu.log
tail -n 30 gpurun_out/pytest_gpu.log
timeout 600 python tools/tune.py --configs c2_flash,c3tp8_flash,c5b16_flash,c1_flash,one_flash,short_flash,c2 --versions v1 > gpurun_out/tune_flash.log 2>&1; echo "exit=$?" >> gpurun_out/tune_flash.log
cat gpurun_out/tune_flash.log
timeout 600 python tools/bench_cache_ops.py > gpurun_out/cache_ops.log 2>&1; echo "exit=$?" >> gpurun_out/cache_ops.log
cat gpurun_out/cache_ops.log
