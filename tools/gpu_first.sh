#!/bin/bash
# usage: tools/gpu_first.sh <tag>  — parity tests, per-kernel breakdown, short bench, then ncu --set full (with source) of one 8M wave
tag=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
BF_WAVE_LANES=8388608 BF_WAVE_STREAMS=1 timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
    --launch-skip 20 --launch-count 20 python tools/kernel_breakdown.py 8388608 > gpurun_out/ncu_full_$tag.log 2>&1
tail -4 gpurun_out/pytest_gpu_$tag.log
cat gpurun_out/breakdown_8M_$tag.log
cat gpurun_out/bench_$tag.json
tail -3 gpurun_out/bench_$tag.err
ls -la gpurun_out/
