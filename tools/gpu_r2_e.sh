#!/bin/bash
# GPU-box script (round 2): the no-alloc entry point's test alone in a fresh process, then the whole -m gpu suite
# (no -x: list every failure; tests/test_zz_no_alloc.py is collected last), smoke.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_no_alloc_alone.log
tail -n 30 gpurun_out/pytest_no_alloc_alone.log
( time timeout 2400 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_gpu.log
tail -n 60 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
