#!/bin/bash
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 120 python tools/simt_probe.py block8 > gpurun_out/simt_probe.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_simt -s 3 -c 1 -o gpurun_out/prof_simt python tools/simt_probe.py block8 > gpurun_out/ncu_simt.log 2>&1
ncu -i gpurun_out/prof_simt.ncu-rep --page raw --csv > gpurun_out/prof_simt_raw.csv 2>/dev/null
tail -3 gpurun_out/ncu_simt.log
