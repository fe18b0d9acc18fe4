#!/bin/bash
# usage: tools/gpu_prof4.sh <tag> <kernel regex> [skip] [count] — ncu --set full (with source) of selected launches of one 16M-sample C3 wave
tag=$1; kern=$2; skip=${3:-1}; cnt=${4:-1}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag -k regex:$kern --launch-skip $skip --launch-count $cnt \
    python tools/c3_wave.py 16777216 > gpurun_out/ncu_full_$tag.log 2>&1
tail -16 gpurun_out/ncu_full_$tag.log
