#!/usr/bin/env bash
# usage: tools/gpu_final.sh TAG : the round's final measurement pass -- GPU tests, smoke, inference + training bench,
# ncu launch list of the bench command, ncu --set full of the CTA-pair linear kernel (enc.s3 fc1 + GELU)
set -x
mkdir -p gpurun_out
tag=${1:-x}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/pytest_$tag.log
tail -3 gpurun_out/pytest_$tag.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3 > gpurun_out/smoke_$tag.log; cat gpurun_out/smoke_$tag.log
timeout 600 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 600 gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err
timeout 600 python bench.py --mode train --no-cpu-baseline > gpurun_out/train_$tag.json 2> gpurun_out/train_$tag.err
tail -c 300 gpurun_out/train_$tag.json; tail -3 gpurun_out/train_$tag.err
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_$tag.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_$tag.log 2>&1
timeout 300 python tools/linear_microbench.py enc.s3 > gpurun_out/plain_lin_$tag.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_linear_kernel -s 30 -c 2 -f -o gpurun_out/prof_linear_pair_$tag \
    python tools/linear_microbench.py enc.s3 > gpurun_out/ncu_lin_$tag.log 2>&1
tail -2 gpurun_out/ncu_lin_$tag.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err
tail -c 400 gpurun_out/bench_ref_$tag.json
