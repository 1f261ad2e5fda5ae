#!/usr/bin/env bash
# round-2 call B: forward-kernel experiment: parity tests of the op, microbench, then a trace build + timeline
set -x
mkdir -p gpurun_out
tag=${1:-r2b}
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "forward or full_size_properties or graphed_inference or blocks_bf16" 2>&1 | tail -8 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
timeout 600 python tools/attn_microbench.py --impl auto > gpurun_out/micro_fwd_$tag.jsonl 2> gpurun_out/micro_fwd_$tag.err
cat gpurun_out/micro_fwd_$tag.jsonl | cut -c1-200
SWIN_PDL=0 timeout 600 python tools/attn_microbench.py --impl auto --only enc_s2 > gpurun_out/micro_fwd_nopdl_$tag.jsonl 2>&1
cat gpurun_out/micro_fwd_nopdl_$tag.jsonl | cut -c1-200
SWIN_TC_TRACE=1 bash monocular-depth-estimation_b200/csrc/build.sh > gpurun_out/build_trace_$tag.log 2>&1
for st in 2 0; do for sh in 0 3; do timeout 120 python tools/tc_trace.py $sh $st > gpurun_out/trace_${tag}_s${st}_sh${sh}.txt 2>&1; done; done
head -50 gpurun_out/trace_${tag}_s2_sh0.txt
