#!/usr/bin/env bash
# usage: tools/gpu_bench.sh TAG  -> gpurun_out/{pytest,bench}_TAG.*
set -x
mkdir -p gpurun_out
tag=${1:-x}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_$tag.log
timeout 600 python bench.py ${BENCH_ARGS:-} > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -4 gpurun_out/pytest_$tag.log
tail -c 2500 gpurun_out/bench_$tag.json
tail -3 gpurun_out/bench_$tag.err
