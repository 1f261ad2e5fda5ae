#!/usr/bin/env bash
set -uo pipefail
tag="${1:-r2unpad12}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$tag.log
timeout 900 python bench.py --version large12 --mode train --no-cpu-baseline > gpurun_out/bench_train_large12_$tag.json 2> gpurun_out/bench_train_large12_$tag.err
echo "bench rc=$?"; python - <<PY
import json
b=json.loads(open('gpurun_out/bench_train_large12_$tag.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ['value','ms_per_step']}, json.dumps(b.get('graph_replay'))[:160], b['clocks'])
r=b.get('roofline'); print('bwd', r['frac'], r['kernel_ms_per_step'], {k:(round(v['us_per_launch'],1),round(v['frac'],3)) for k,v in r['by_shape'].items()})
PY
tail -2 gpurun_out/bench_train_large12_$tag.err
