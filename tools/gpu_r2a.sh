#!/usr/bin/env bash
# round-2 call A: full GPU test-suite (all failures listed), the default bench line, the reference arm
set -x
mkdir -p gpurun_out
tag=${1:-r2a}
rm -f gpurun_out/parity_errors.jsonl
timeout 1200 python -m pytest tests -m gpu -q -rf 2>&1 | tail -60 > gpurun_out/pytest_$tag.log
tail -30 gpurun_out/pytest_$tag.log
cp gpurun_out/parity_errors.jsonl gpurun_out/parity_errors_$tag.jsonl
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 3000 gpurun_out/bench_$tag.json; tail -5 gpurun_out/bench_$tag.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err
tail -c 1500 gpurun_out/bench_ref_$tag.json; tail -3 gpurun_out/bench_ref_$tag.err
