#!/usr/bin/env bash
# full validation of a forward-kernel build: GPU suite (-x), metric diag, forward microbench, default bench, ncu --set full (stage 2 and 0)
set -x
mkdir -p gpurun_out
tag=${1:-r2k}
rm -f gpurun_out/parity_errors.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q -rf 2>&1 | tail -12 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
cp gpurun_out/parity_errors.jsonl gpurun_out/parity_errors_$tag.jsonl
python tools/diag_metric_noise.py > gpurun_out/diag_metric_$tag.jsonl 2>/dev/null; cut -c1-330 gpurun_out/diag_metric_$tag.jsonl
timeout 600 python tools/attn_microbench.py --impl auto > gpurun_out/micro_fwd_$tag.jsonl 2> gpurun_out/micro_fwd_$tag.err
cut -c1-150 gpurun_out/micro_fwd_$tag.jsonl
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 600 gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err
bash tools/gpu_ncu_full.sh ${tag}_fwd_s2 enc_s2_352
bash tools/gpu_ncu_full.sh ${tag}_fwd_s0 enc_s0_352
