#!/usr/bin/env bash
# full GPU suite (-x) + default bench line
set -x
mkdir -p gpurun_out
tag=${1:-r2t}
rm -f gpurun_out/parity_errors.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q -rf 2>&1 | tail -12 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
cp gpurun_out/parity_errors.jsonl gpurun_out/parity_errors_$tag.jsonl
timeout 1200 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 400 gpurun_out/bench_$tag.json; tail -5 gpurun_out/bench_$tag.err
