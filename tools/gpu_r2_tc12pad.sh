#!/usr/bin/env bash
# un-aligned window-12 forward: parity tests, then the microbench rows
set -uo pipefail
tag="${1:-r2tc12pad}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -k "ws12 or forward_matches_oracle" > gpurun_out/pytest_$tag.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$tag.log
timeout 300 python tools/attn_microbench.py --only oda_ws12 > gpurun_out/micro_$tag.jsonl 2> gpurun_out/micro_$tag.err
echo "micro rc=$?"; cat gpurun_out/micro_$tag.jsonl
timeout 600 python -m pytest tests/test_siblings_gpu.py -x -q > gpurun_out/pytest_sib_$tag.log 2>&1
echo "siblings rc=$?"; tail -3 gpurun_out/pytest_sib_$tag.log
