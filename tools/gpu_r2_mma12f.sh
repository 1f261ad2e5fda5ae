#!/usr/bin/env bash
set -uo pipefail
tag="${1:-r2mma12f}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -k "ws12 or forward_matches_oracle or backward_matches_oracle" > gpurun_out/pytest_$tag.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/pytest_$tag.log
timeout 300 python tools/attn_microbench.py --only oda_ws12 --impl auto > gpurun_out/micro_$tag.jsonl 2> gpurun_out/micro_$tag.err
echo "micro rc=$?"; cat gpurun_out/micro_$tag.jsonl; tail -3 gpurun_out/micro_$tag.err
timeout 300 python tools/attn_microbench.py --only oda_ws12 --impl mma > gpurun_out/micro_${tag}_mma.jsonl 2> gpurun_out/micro_${tag}_mma.err
cat gpurun_out/micro_${tag}_mma.jsonl
timeout 600 python -m pytest tests/test_siblings_gpu.py -x -q > gpurun_out/pytest_sib_$tag.log 2>&1
echo "siblings rc=$?"; tail -3 gpurun_out/pytest_sib_$tag.log
