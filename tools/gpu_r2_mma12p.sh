#!/usr/bin/env bash
set -uo pipefail
tag="${1:-r2mma12p}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_gpu.py -x -q -k "ws12 or window12 or backward_matches_oracle or guard" > gpurun_out/pytest_$tag.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$tag.log
timeout 300 python tools/attn_microbench.py --only oda_ws12 --impl mma > gpurun_out/micro_${tag}_fwd.jsonl 2> gpurun_out/micro_${tag}.err
timeout 300 python tools/attn_microbench.py --only oda_ws12 --impl auto --bwd > gpurun_out/micro_${tag}_bwd.jsonl 2>> gpurun_out/micro_${tag}.err
cat gpurun_out/micro_${tag}_fwd.jsonl gpurun_out/micro_${tag}_bwd.jsonl | cut -c1-150
