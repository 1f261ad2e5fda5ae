#!/usr/bin/env bash
set -uo pipefail
tag="${1:-r2train}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_reference_gpu.py -x -q -k "trainer_clip or drop_path_gates or graphed_train or training or nyu" > gpurun_out/pytest_$tag.log 2>&1
echo "pytest rc=$?"; tail -6 gpurun_out/pytest_$tag.log
timeout 900 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_$tag.json 2> gpurun_out/bench_train_$tag.err
echo "bench rc=$?"; python - <<PY
import json
b=json.loads(open('gpurun_out/bench_train_$tag.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ['value','ms_per_step','gpu_launches']}, b.get('eager'), b.get('e2e',{}).get('value'))
PY
tail -2 gpurun_out/bench_train_$tag.err
