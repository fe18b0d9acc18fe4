#!/bin/bash
# usage: tools/gpu_tests.sh <tag> [pytest args] — GPU parity tests only
tag=${1:-x}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -q "$@" > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
grep -E "^(FAILED|ERROR)|passed|failed|^E  " gpurun_out/pytest_gpu_$tag.log | tail -40
