#!/bin/bash
# GPU-box script: the newest parity suites (rope, cache ops incl. flash layout), smoke, one bench line.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_rope.py tests/test_gpu_cache_ops.py -m gpu -x -q ) > gpurun_out/pytest_new.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_new.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
tail -n 5 gpurun_out/pytest_new.log gpurun_out/smoke.log gpurun_out/bench_ours.json
