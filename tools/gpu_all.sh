#!/usr/bin/env bash
# usage: tools/gpu_all.sh TAG : GPU tests, bench, ncu launch list of the bench, forward microbench
set -x
mkdir -p gpurun_out
tag=${1:-x}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_$tag.log
tail -4 gpurun_out/pytest_$tag.log
timeout 600 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 2500 gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_$tag.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_$tag.log 2>&1
timeout 600 python tools/attn_microbench.py --impl auto ${MICRO_ARGS:-} > gpurun_out/micro_$tag.jsonl 2> gpurun_out/micro_$tag.err
cat gpurun_out/micro_$tag.jsonl
