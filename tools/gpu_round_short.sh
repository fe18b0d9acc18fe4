#!/bin/bash
# usage: tools/gpu_round_short.sh <tag> — gpu_round.sh without the parity suite: breakdown, bench (both arms), ncu launch list, ncu --set full of one wave
tag=${1:-x}
mkdir -p gpurun_out
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --samples 67108864 --no-cpu-baseline --no-extras > gpurun_out/ncu_launches_$tag.log 2>&1
BF_WAVE_LANES=8388608 BF_WAVE_STREAMS=1 timeout 400 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
    --launch-skip 17 --launch-count 17 python tools/kernel_breakdown.py 8388608 > gpurun_out/ncu_full_$tag.log 2>&1
cat gpurun_out/breakdown_8M_$tag.log | head -3
python -c "
import json; d = json.load(open('gpurun_out/bench_$tag.json')); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'], d['clocks'])"
