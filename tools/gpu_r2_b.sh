#!/bin/bash
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_attention.py tests/test_gpu_properties.py -m gpu -x -q ) > gpurun_out/pytest_attn.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_attn.log
tail -n 6 gpurun_out/pytest_attn.log
( time timeout 2400 python -m pytest tests/test_gpu_reference_suite.py -m gpu -q ) > gpurun_out/pytest_refsuite.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_refsuite.log
tail -n 40 gpurun_out/pytest_refsuite.log
timeout 600 python tests/golden/make_golden_cuda.py gpurun_out/ref_cuda_golden.npz > gpurun_out/make_golden_cuda.log 2>&1; echo "exit=$?" >> gpurun_out/make_golden_cuda.log
tail -n 25 gpurun_out/make_golden_cuda.log
timeout 1200 python tools/tune.py --configs c2,c3tp8,c3tp4,c3tp2,c1,c4,c5b16,c5b32,small,one,mid,short --versions v1,v2 --knobs "" "split_cost=256" "split_cost=1024" "stages=3" > gpurun_out/tune_grid_fused.log 2>&1; echo "exit=$?" >> gpurun_out/tune_grid_fused.log
cat gpurun_out/tune_grid_fused.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
tail -n 5 gpurun_out/bench_ours.err; cat gpurun_out/bench_ours.json
