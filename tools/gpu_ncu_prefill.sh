#!/bin/bash
# GPU-box script: ncu full capture of the prefill kernel (after a clean plain run).
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
CMD="python tools/bench_prefill.py --cases llama8b_chunk2k_ctx4k"
timeout 300 $CMD > gpurun_out/prefill_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:context_attention -s 3 -c 1 -o gpurun_out/prof_prefill $CMD > gpurun_out/ncu_prefill.log 2>&1
ncu -i gpurun_out/prof_prefill.ncu-rep --page raw --csv > gpurun_out/prof_prefill_raw.csv 2>/dev/null
ls -la gpurun_out
