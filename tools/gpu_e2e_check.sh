#!/usr/bin/env bash
# serving-loop tests + the inference bench line twice: e2e and its five regions
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -q -k "infer_host_batches or graphed_inference" 2>&1 | tail -3
for i in 1 2; do
  python bench.py --no-cpu-baseline --no-reference-gpu --no-train-leg --no-graph-timing 2>/dev/null | python -c '
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d["value"], 1), round(d["e2e"]["value"], 1), [round(x, 1) for x in d["e2e"]["regions_images_per_s"]],
      round(d["e2e_eager"]["value"], 1), round(d["graph_replay"]["value"], 1))'
done
