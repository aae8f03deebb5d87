#!/bin/bash
# GPU-box script (round 2): parity of the decode kernel with the fused split-K reduce + timing on every config.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_attention.py tests/test_gpu_properties.py -m gpu -x -q ) > gpurun_out/pytest_attn.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_attn.log
tail -n 15 gpurun_out/pytest_attn.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
timeout 1200 python tools/tune.py --configs c2,c3tp8,c3tp4,c3tp2,c1,c4,c5b16,c5b32,small,one,mid,short --versions v1,v2 --knobs "" "ctas_per_sm_target=4" "ctas_per_sm_target=16" "stages=2" > gpurun_out/tune_ab.log 2>&1; echo "exit=$?" >> gpurun_out/tune_ab.log
cat gpurun_out/tune_ab.log
