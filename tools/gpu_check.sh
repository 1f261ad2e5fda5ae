#!/usr/bin/env bash
# usage: tools/gpu_check.sh TAG : GPU tests, smoke, inference bench, training bench (no ncu)
set -x
mkdir -p gpurun_out
tag=${1:-x}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_$tag.log
tail -4 gpurun_out/pytest_$tag.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4 > gpurun_out/smoke_$tag.log; cat gpurun_out/smoke_$tag.log
timeout 600 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 2500 gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err
timeout 600 python bench.py --mode train --no-cpu-baseline > gpurun_out/train_$tag.json 2> gpurun_out/train_$tag.err
tail -c 2500 gpurun_out/train_$tag.json; tail -3 gpurun_out/train_$tag.err
