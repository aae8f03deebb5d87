#!/bin/bash
# GPU-box script: properties tests, smoke, bench (both arms), ncu launch list + full capture of the top kernel.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_properties.py -m gpu -x -q > gpurun_out/pytest_props.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_props.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "exit=$?" >> gpurun_out/bench_ours.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
nproc > gpurun_out/host.txt; lscpu | head -20 >> gpurun_out/host.txt; nvidia-smi >> gpurun_out/host.txt
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:paged_attention_mma -s 5 -c 2 -o gpurun_out/prof_mma $CMD > gpurun_out/ncu_full.log 2>&1
tail -n 3 gpurun_out/pytest_props.log gpurun_out/smoke.log gpurun_out/bench_ours.json gpurun_out/bench_ref.json
