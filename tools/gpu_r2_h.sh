#!/bin/bash
# GPU-box script (round 2): parity of the decode kernels after the SIMT inner-loop change + the bench's SIMT lines.
set -x
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_attention.py tests/test_gpu_properties.py tests/test_golden_cuda.py -m gpu -x -q ) > gpurun_out/pytest_attn.log 2>&1; echo "exit=$?" >> gpurun_out/pytest_attn.log
tail -n 8 gpurun_out/pytest_attn.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "exit=$?" >> gpurun_out/bench_short.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_short.json").read())
for c in d["extra"]["configs"]:
    if c["name"].startswith("simt"):
        print(c["name"], c["ms"], "ms", c["gbps"], "GB/s", c["kernel"])
print("headline", d["value"], "prefill", d["extra"]["prefill"]["algorithmic_tflops"])
PY
