#!/usr/bin/env bash
# usage: tools/gpu_scale.sh N TAG : inference and training bench at N GPUs of one box (torchrun, as the driver launches it)
set -x
n=${1:-2}; tag=${2:-x}
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench${n}_$tag.json 2> gpurun_out/bench${n}_$tag.err
tail -2 gpurun_out/bench${n}_$tag.err
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $n --steps 10 --warmup 3 --mode train > gpurun_out/train${n}_$tag.json 2> gpurun_out/train${n}_$tag.err
tail -2 gpurun_out/train${n}_$tag.err
