#!/bin/bash
# GPU-box script (round 2): block-sparse decode on the mma family — parity (both block-sparse tests) and timing.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_attention.py -m gpu -x -q -k blocksparse ) > gpurun_out/pytest_blocksparse.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_blocksparse.log
tail -n 25 gpurun_out/pytest_blocksparse.log
timeout 300 python tools/simt_probe.py sweep > gpurun_out/simt_sweep.log 2>&1; cat gpurun_out/simt_sweep.log
timeout 300 python tools/gpu_quickcheck.py --time --no-check 2>&1 | grep -i "sparse\|block8\|fp32" > gpurun_out/quickcheck_sparse.log; cat gpurun_out/quickcheck_sparse.log
