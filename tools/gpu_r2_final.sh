#!/usr/bin/env bash
# usage: tools/gpu_r2_final.sh TAG : round-2 measurement pass -- GPU suite, smoke, default bench line, reference arm, ncu
# launch list of the bench command, ncu --set full of the backward kernel (stage 0 of the benchmarked grid, stage 2 of the
# training grid), forward / backward / global-attention microbench sweeps
set -x
mkdir -p gpurun_out
tag=${1:-r2f}
rm -f gpurun_out/parity_errors.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/pytest_$tag.log
tail -3 gpurun_out/pytest_$tag.log
cp gpurun_out/parity_errors.jsonl gpurun_out/parity_errors_$tag.jsonl
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3 > gpurun_out/smoke_$tag.log; cat gpurun_out/smoke_$tag.log
timeout 1200 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 300 gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err
tail -c 400 gpurun_out/bench_ref_$tag.json
NCU_BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-reference-gpu --no-train-leg --no-graph-timing"
timeout 300 $NCU_BENCH > gpurun_out/plain_$tag.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_$tag.csv \
    $NCU_BENCH > gpurun_out/ncu_launches_$tag.log 2>&1
KREGEX=tc_bwd bash tools/gpu_ncu_full.sh ${tag}_bwd_s0 enc_s0_352 --bwd
KREGEX=tc_bwd bash tools/gpu_ncu_full.sh ${tag}_bwd_s2 enc_s2_480x640 --bwd
timeout 600 python tools/attn_microbench.py --impl auto > gpurun_out/micro_fwd_$tag.jsonl 2> gpurun_out/micro_fwd_$tag.err
timeout 600 python tools/attn_microbench.py --impl auto --bwd > gpurun_out/micro_bwd_$tag.jsonl 2> gpurun_out/micro_bwd_$tag.err
python tools/mha_microbench.py > gpurun_out/micro_mha_$tag.jsonl 2>&1
tail -3 gpurun_out/micro_mha_$tag.jsonl
