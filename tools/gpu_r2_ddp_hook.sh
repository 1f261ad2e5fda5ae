#!/usr/bin/env bash
# usage: tools/gpu_r2_ddp_hook.sh N TAG : the DDP training step at N GPUs, captured in a CUDA graph, with fp32 gradient
# buckets (default) and with DDP's bf16_compress_hook (--bf16-grad-hook)
set -x
mkdir -p gpurun_out
n=${1:-2}; tag=${2:-r2}
for hook in "" "--bf16-grad-hook"; do
  name=$([ -z "$hook" ] && echo fp32 || echo bf16hook)
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $n --mode train --steps 10 --warmup 3 --no-cpu-baseline --no-graph-timing $hook \
      > gpurun_out/train_${n}gpu_${name}_$tag.json 2> gpurun_out/train_${n}gpu_${name}_$tag.err
  echo "rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/train_${n}gpu_${name}_$tag.json").read().strip().splitlines()[-1])
print("$name", {k: d.get(k) for k in ("value", "ms_per_step", "n_gpus")}, {k: d["config"].get(k) for k in ("graph_replay", "ddp_replicas_in_sync", "bf16_grad_hook")} if "config" in d else "", d.get("eager"))
PY
done
