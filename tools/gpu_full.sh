#!/bin/bash
# usage: tools/gpu_full.sh <tag> — ncu --set full (with source) of the 17 launches of one 8M-sample C2 wave (the second wave of kernel_breakdown.py)
tag=${1:-x}
mkdir -p gpurun_out
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1 || exit 1
BF_WAVE_LANES=8388608 BF_WAVE_STREAMS=1 timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
    --launch-skip 17 --launch-count 17 python tools/kernel_breakdown.py 8388608 > gpurun_out/ncu_full_$tag.log 2>&1
tail -3 gpurun_out/ncu_full_$tag.log
