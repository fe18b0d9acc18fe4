#!/bin/bash
# usage: tools/gpu_prof2.sh <tag> <lanes> <kernel regex> <skip> <count>  — ncu --set full of selected launches of one big wave
tag=$1; lanes=$2; kern=$3; skip=$4; cnt=$5
mkdir -p gpurun_out
python tools/kernel_breakdown.py $lanes > gpurun_out/breakdown_${lanes}_$tag.log 2>&1
cat gpurun_out/breakdown_${lanes}_$tag.log
BF_WAVE_STREAMS=1 timeout 1200 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
    -k regex:$kern --launch-skip $skip --launch-count $cnt python tools/kernel_breakdown.py $lanes > gpurun_out/ncu_full_$tag.log 2>&1
tail -3 gpurun_out/ncu_full_$tag.log
