#!/bin/bash
# GPU-box script (round 2): measurement artefacts on the product build — both bench arms, ncu launch list of the bench,
# full ncu captures (decode kernel on config 2 inside the bench command, prefill kernel), prefill timings.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 900 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
tail -n 3 gpurun_out/bench_ours.err; cat gpurun_out/bench_ours.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_mma -s 5 -c 1 -o gpurun_out/prof_mma $CMD > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/prof_mma.ncu-rep --page raw --csv > gpurun_out/prof_mma_raw.csv 2>/dev/null
PCMD="python tools/bench_prefill.py --cases llama8b_chunk2k_ctx4k"
timeout 300 $PCMD > gpurun_out/prefill_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:context_attention -s 3 -c 1 -o gpurun_out/prof_prefill $PCMD > gpurun_out/ncu_prefill.log 2>&1
ncu -i gpurun_out/prof_prefill.ncu-rep --page raw --csv > gpurun_out/prof_prefill_raw.csv 2>/dev/null
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill.log
cat gpurun_out/bench_prefill.log
