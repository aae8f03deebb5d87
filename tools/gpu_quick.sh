#!/bin/bash
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_rope.py -m gpu -q ) > gpurun_out/pytest_rope.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_rope.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
tail -n 4 gpurun_out/pytest_rope.log gpurun_out/smoke.log; cat gpurun_out/bench_ours.json; tail -n 3 gpurun_out/bench_ours.err
