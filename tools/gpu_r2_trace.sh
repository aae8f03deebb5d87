#!/bin/bash
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python tools/sk_trace.py c2 c3tp8 c5b16 > gpurun_out/sk_trace.log 2>&1; echo "exit=$?" >> gpurun_out/sk_trace.log
cat gpurun_out/sk_trace.log
timeout 900 python tools/tune.py --configs c2,c3tp8,c3tp4,c5b16,mid,one,c1 --versions v2 --knobs "" "grid_ctas=592" "grid_ctas=1184" "grid_ctas=2368" "grid_ctas=4736" "grid_kernel=1" > gpurun_out/tune_grid.log 2>&1; echo "exit=$?" >> gpurun_out/tune_grid.log
cat gpurun_out/tune_grid.log
