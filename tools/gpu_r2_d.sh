#!/bin/bash
# GPU-box script (round 2, re-entry): whole -m gpu suite on the current tree, smoke, both bench arms, ncu launch list of
# the bench, one full ncu capture of the decode kernel on the config-3 TP8 shard, flash-layout + cache-op timings.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/host.txt 2>&1
( time timeout 2000 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_gpu.log
tail -n 8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
tail -n 3 gpurun_out/bench_ours.err; cat gpurun_out/bench_ours.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
TCMD="python tools/tune.py --configs c3tp8 --versions v2"
timeout 300 $TCMD > gpurun_out/tune_c3tp8.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_mma -s 3 -c 1 -o gpurun_out/prof_c3tp8 $TCMD > gpurun_out/ncu_c3tp8.log 2>&1
ncu -i gpurun_out/prof_c3tp8.ncu-rep --page raw --csv > gpurun_out/prof_c3tp8_raw.csv 2>/dev/null
timeout 600 python tools/tune.py --configs c1,c2,c3tp2,c3tp4,c3tp8,c4,c5b16,c5b32,small,one,mid,short --versions v1,v2 > gpurun_out/tune_all.log 2>&1; echo "exit=$?" >> gpurun_out/tune_all.log
cat gpurun_out/tune_all.log
timeout 600 python tools/tune.py --configs c2_flash,c3tp8_flash,c5b16_flash,c1_flash,one_flash,short_flash --versions v1 > gpurun_out/tune_flash.log 2>&1; echo "exit=$?" >> gpurun_out/tune_flash.log
cat gpurun_out/tune_flash.log
timeout 600 python tools/bench_cache_ops.py > gpurun_out/cache_ops.log 2>&1; echo "exit=$?" >> gpurun_out/cache_ops.log
cat gpurun_out/cache_ops.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1; echo "exit=$?" >> gpurun_out/bench_prefill.log
cat gpurun_out/bench_prefill.log
