#!/usr/bin/env bash
# run one test N times, report failures with their assertion output
mkdir -p gpurun_out; : > gpurun_out/flake.log
n=${2:-12}; fails=0
for i in $(seq 1 $n); do
  timeout 300 python -m pytest tests/test_parity_gpu.py -x -q -k "$1" > gpurun_out/flake_one.log 2>&1 || { fails=$((fails+1)); grep -n "^E \|assert\|Error" gpurun_out/flake_one.log | head -12 >> gpurun_out/flake.log; echo "---- run $i" >> gpurun_out/flake.log; }
done
echo "fails=$fails of $n"; cat gpurun_out/flake.log | head -60
