#!/bin/bash
# GPU-box script: whole `-m gpu` suite, smoke, bench (both arms), ncu launch list + full captures (decode mma kernel,
# tcgen05 prefill kernel), timing sweeps.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_gpu.log
( B200PA_PREFILL_WS=0 timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_fused.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_fused.log
( B200PA_PREFILL_TC=0 timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_mma.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_mma.log
for nq in 1 3; do ( B200PA_PREFILL_NQ=$nq timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_nq$nq.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_nq$nq.log; done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:paged_attention_mma -s 5 -c 1 -o gpurun_out/prof_mma $CMD > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/prof_mma.ncu-rep --page raw --csv > gpurun_out/prof_mma_raw.csv 2>/dev/null
PCMD="python tools/bench_prefill.py --cases llama8b_chunk2k_ctx4k"
timeout 300 $PCMD > gpurun_out/prefill_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:context_attention -s 3 -c 1 -o gpurun_out/prof_prefill $PCMD > gpurun_out/ncu_prefill.log 2>&1
ncu -i gpurun_out/prof_prefill.ncu-rep --page raw --csv > gpurun_out/prof_prefill_raw.csv 2>/dev/null
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1
timeout 600 python tools/tune.py --configs c1,c2,c3tp2,c3tp8,c4,c5b16,c5b32,small,one --versions v1,v2 > gpurun_out/tune_all.log 2>&1
timeout 300 python tools/bench_cache_ops.py > gpurun_out/cache_ops.log 2>&1
tail -n 5 gpurun_out/pytest_gpu.log gpurun_out/pytest_prefill_fused.log gpurun_out/pytest_prefill_mma.log gpurun_out/pytest_prefill_nq1.log gpurun_out/pytest_prefill_nq3.log gpurun_out/smoke.log gpurun_out/bench_ours.json gpurun_out/bench_ref.json
