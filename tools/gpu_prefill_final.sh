#!/bin/bash
# GPU-box script: prefill kernels only — probe, parity suite in every Q-tile pairing mode and kernel family, smoke,
# bench line, timing vs the reference Triton kernels, one ncu --set full capture of the tcgen05 kernel.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
bash tools/gpu_prefill.sh > gpurun_out/prefill_default.log 2>&1
for nq in 1 3; do ( B200PA_PREFILL_NQ=$nq timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_nq$nq.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_nq$nq.log; done
( B200PA_PREFILL_WS=0 timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_fused.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_fused.log
( B200PA_PREFILL_TC=0 timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_mma.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_mma.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
PCMD="python tools/bench_prefill.py --cases llama8b_chunk2k_ctx4k"
timeout 300 $PCMD > gpurun_out/prefill_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:context_attention -s 3 -c 1 -o gpurun_out/prof_prefill $PCMD > gpurun_out/ncu_prefill.log 2>&1
ncu -i gpurun_out/prof_prefill.ncu-rep --page raw --csv > gpurun_out/prof_prefill_raw.csv 2>/dev/null
tail -n 14 gpurun_out/prefill_default.log | cut -c1-200
tail -n 3 gpurun_out/pytest_prefill_nq1.log gpurun_out/pytest_prefill_nq3.log gpurun_out/pytest_prefill_fused.log gpurun_out/pytest_prefill_mma.log gpurun_out/smoke.log
cut -c1-1500 gpurun_out/bench_ours.json
