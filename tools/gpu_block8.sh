#!/bin/bash
# GPU-box script: decode-kernel validation after the block-size-8 variant — the decode-related parity files (-x, as the
# driver runs them) + timing against the reference's CUDA kernels.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_attention.py tests/test_gpu_properties.py tests/test_golden_cuda.py tests/test_gpu_flash_decode.py tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/pytest_decode.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_decode.log
tail -n 12 gpurun_out/pytest_decode.log
timeout 300 python tools/gpu_quickcheck.py --time --no-check > gpurun_out/quickcheck_vs_reference_cuda.log 2>&1; cat gpurun_out/quickcheck_vs_reference_cuda.log
