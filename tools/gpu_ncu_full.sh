#!/usr/bin/env bash
# usage: tools/gpu_ncu_full.sh TAG CASE [--bwd] : ncu --set full of the tcgen05 kernel on one microbench case
set -x
mkdir -p gpurun_out
tag=$1; case=$2; extra=${3:-}
timeout 300 python tools/attn_microbench.py --only $case --iters 2 --impl auto $extra > gpurun_out/plain_$tag.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-tc_} -s 8 -c 2 -f -o gpurun_out/prof_$tag \
    python tools/attn_microbench.py --only $case --iters 2 --impl auto $extra > gpurun_out/ncu_$tag.log 2>&1
tail -2 gpurun_out/ncu_$tag.log
