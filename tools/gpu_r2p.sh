#!/usr/bin/env bash
# backward-kernel iteration: backward parity tests, backward microbench (product build), then a trace build on the box and the timelines
set -x
mkdir -p gpurun_out
tag=${1:-r2p}
rm -f gpurun_out/parity_errors.jsonl
timeout 200 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "bwd or backward" 2>&1 | tail -3; timeout 1200 python -m pytest tests -m gpu -x -q -rf -k "bwd or backward or grad or train or unpadded or no_alloc" 2>&1 | tail -12 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
timeout 600 python tools/attn_microbench.py --impl auto --bwd > gpurun_out/micro_bwd_$tag.jsonl 2> gpurun_out/micro_bwd_$tag.err
cut -c1-170 gpurun_out/micro_bwd_$tag.jsonl
cp -r monocular-depth-estimation_b200 /tmp/pkg_backup
SWIN_TC_TRACE=1 bash monocular-depth-estimation_b200/csrc/build.sh > gpurun_out/trace_build_$tag.log 2>&1
for st in 2 0 3; do python tools/tc_trace_bwd.py $st 0 > gpurun_out/trace_bwd_${tag}_s${st}.txt 2>&1; done
head -8 gpurun_out/trace_bwd_${tag}_s2.txt | cut -c1-200
