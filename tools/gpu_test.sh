#!/usr/bin/env bash
# usage: tools/gpu_test.sh TAG [pytest -k expression]
mkdir -p gpurun_out
tag=${1:-x}
timeout ${TEST_TIMEOUT:-600} python -m pytest tests -m gpu -x -q ${2:+-k "$2"} 2>&1 | tail -40 > gpurun_out/pytest_$tag.log
cat gpurun_out/pytest_$tag.log
