#!/bin/bash
# GPU-box script (round 2): prefill kernel with TMA staging (parity + timing, against the cp.async loaders), the
# reference-CUDA comparison tests (ref_cuda.so present), and the round's measurement artefacts: both bench arms, ncu
# launch list of the bench, full ncu captures (decode kernel on the config-3 TP8 shard, prefill kernel), sweeps.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total,driver_version --format=csv > gpurun_out/host.txt 2>&1
( time timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q ) > gpurun_out/pytest_prefill_tma.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_prefill_tma.log
tail -n 12 gpurun_out/pytest_prefill_tma.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill_tma.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill_tma.log
cat gpurun_out/bench_prefill_tma.log
B200PA_PREFILL_TMA=0 timeout 600 python tools/bench_prefill.py --cases llama8b_chunk2k_ctx4k,llama8b_chunk512_ctx8k,llama8b_prefix1k_new1k,llama70b_tp8_chunk2k_ctx4k,llama7b_mha_chunk2k_ctx4k > gpurun_out/bench_prefill_cpasync.log 2>&1
cat gpurun_out/bench_prefill_cpasync.log
( time timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cache_ops.py tests/test_gpu_rope.py -m gpu -x -q -k "reference_cuda or ref_cuda" ) > gpurun_out/pytest_vs_ref_cuda.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_vs_ref_cuda.log
tail -n 6 gpurun_out/pytest_vs_ref_cuda.log
timeout 900 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
tail -n 3 gpurun_out/bench_ours.err; cat gpurun_out/bench_ours.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
TCMD="python tools/tune.py --configs c3tp8 --versions v2"
timeout 300 $TCMD > gpurun_out/tune_c3tp8.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_mma -s 3 -c 1 -o gpurun_out/prof_c3tp8 $TCMD > gpurun_out/ncu_c3tp8.log 2>&1
ncu -i gpurun_out/prof_c3tp8.ncu-rep --page raw --csv > gpurun_out/prof_c3tp8_raw.csv 2>/dev/null
PCMD="python tools/bench_prefill.py --cases llama8b_chunk2k_ctx4k"
timeout 300 $PCMD > gpurun_out/prefill_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:context_attention -s 3 -c 1 -o gpurun_out/prof_prefill $PCMD > gpurun_out/ncu_prefill.log 2>&1
ncu -i gpurun_out/prof_prefill.ncu-rep --page raw --csv > gpurun_out/prof_prefill_raw.csv 2>/dev/null
timeout 600 python tools/tune.py --configs c1,c2,c3tp2,c3tp4,c3tp8,c4,c5b16,c5b32,small,one,mid,short --versions v1,v2 > gpurun_out/tune_all.log 2>&1; echo "exit=$?" >> gpurun_out/tune_all.log
cat gpurun_out/tune_all.log
timeout 600 python tools/tune.py --configs c2_flash,c3tp8_flash,c5b16_flash,c1_flash,one_flash,short_flash --versions v1 > gpurun_out/tune_flash.log 2>&1; echo "exit=$?" >> gpurun_out/tune_flash.log
cat gpurun_out/tune_flash.log
timeout 600 python tools/bench_cache_ops.py > gpurun_out/cache_ops.log 2>&1; echo "exit=$?" >> gpurun_out/cache_ops.log
cat gpurun_out/cache_ops.log
