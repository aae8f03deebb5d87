#!/bin/bash
# GPU-box script: the whole `-m gpu` suite, smoke, and a timing sweep over the BASELINE configs.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
timeout 600 python tools/tune.py --configs c1,c2,c3tp2,c3tp8,c4,c5b16,c5b32,small,one --versions v1,v2 > gpurun_out/tune_all.log 2>&1
timeout 300 python tools/bench_cache_ops.py > gpurun_out/cache_ops.log 2>&1
tail -n 5 gpurun_out/pytest_gpu.log gpurun_out/smoke.log
cat gpurun_out/tune_all.log
