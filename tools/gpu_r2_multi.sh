#!/usr/bin/env bash
# usage: tools/gpu_r2_multi.sh N TAG : the default bench line (inference + train leg) at N GPUs, as the driver launches it
set -x
mkdir -p gpurun_out
n=${1:-2}; tag=${2:-r2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_${n}gpu_$tag.json 2> gpurun_out/bench_${n}gpu_$tag.err
echo "rc=$?"
tail -c 1500 gpurun_out/bench_${n}gpu_$tag.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_${n}gpu_$tag.json").read().strip().splitlines()[-1])
t = d.get("train", {})
print({"n": d["n_gpus"], "value": round(d["value"], 1), "e2e": round(d["e2e"]["value"], 1), "frac": round(d["roofline"]["frac"], 3),
       "train": {k: t.get(k) for k in ("value", "ms_per_step", "value_is", "graph_replay", "graph_error", "ddp_replicas_in_sync", "error")},
       "train_eager": t.get("eager")})
PY
