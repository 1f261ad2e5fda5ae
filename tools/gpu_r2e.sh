#!/usr/bin/env bash
# full GPU suite with -x (new no-alloc test last), the no-alloc test alone, metric diag, forward microbench
set -x
mkdir -p gpurun_out
tag=${1:-r2e}
rm -f gpurun_out/parity_errors.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q -rf 2>&1 | tail -12 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
cp gpurun_out/parity_errors.jsonl gpurun_out/parity_errors_$tag.jsonl
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/pytest_noalloc_alone_$tag.log
cat gpurun_out/pytest_noalloc_alone_$tag.log
python tools/diag_metric_noise.py > gpurun_out/diag_metric_$tag.jsonl 2>/dev/null; cat gpurun_out/diag_metric_$tag.jsonl | cut -c1-330
timeout 600 python tools/attn_microbench.py --impl auto > gpurun_out/micro_fwd_$tag.jsonl 2> gpurun_out/micro_fwd_$tag.err
cat gpurun_out/micro_fwd_$tag.jsonl | cut -c1-150
