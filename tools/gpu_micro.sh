#!/usr/bin/env bash
# usage: tools/gpu_micro.sh TAG [ncu-kernel-regex] : microbench + optional ncu --set full of one case
set -x
mkdir -p gpurun_out
tag=${1:-x}
timeout 600 python tools/attn_microbench.py ${MICRO_ARGS:-} > gpurun_out/micro_$tag.jsonl 2> gpurun_out/micro_$tag.err
cat gpurun_out/micro_$tag.jsonl
if [ -n "${2:-}" ]; then
  timeout 300 python tools/attn_microbench.py --only ${NCU_CASE:-enc_s2_352} --iters 2 --impl auto > gpurun_out/plain_micro.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -s 4 -c 2 -o gpurun_out/prof_$tag \
      python tools/attn_microbench.py --only ${NCU_CASE:-enc_s2_352} --iters 2 --impl auto > gpurun_out/ncu_$tag.log 2>&1
  tail -3 gpurun_out/ncu_$tag.log
fi
