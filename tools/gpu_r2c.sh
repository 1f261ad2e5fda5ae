#!/usr/bin/env bash
# forward-kernel experiment loop: op parity tests, microbench, trace build + timeline of stage 2 / 0
set -x
mkdir -p gpurun_out
tag=${1:-r2c}
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "forward or backward or full_size or graphed_inference or blocks_bf16 or crf_block or swin_block" 2>&1 | tail -8 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
timeout 600 python tools/attn_microbench.py --impl auto ${MICRO_ONLY:+--only $MICRO_ONLY} > gpurun_out/micro_fwd_$tag.jsonl 2> gpurun_out/micro_fwd_$tag.err
cat gpurun_out/micro_fwd_$tag.jsonl | cut -c1-200
if [ "${TRACE:-1}" = "1" ]; then
SWIN_TC_TRACE=1 bash monocular-depth-estimation_b200/csrc/build.sh > gpurun_out/build_trace_$tag.log 2>&1
for st in 2 0; do for sh in 0 3; do timeout 120 python tools/tc_trace.py $sh $st > gpurun_out/trace_${tag}_s${st}_sh${sh}.txt 2>&1; done; done
head -70 gpurun_out/trace_${tag}_s2_sh0.txt
head -3 gpurun_out/trace_${tag}_s0_sh0.txt
fi
