#!/bin/bash
# usage: tools/gpu_check.sh <tag>   (run on the GPU box through gpurun)
tag=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python tools/kernel_breakdown.py 1048576 > gpurun_out/breakdown_1M_$tag.log 2>&1
python bench.py > gpurun_out/bench_$tag.log 2> gpurun_out/bench_$tag.err
tail -3 gpurun_out/pytest_gpu_$tag.log
cat gpurun_out/bench_$tag.log
