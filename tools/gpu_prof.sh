#!/bin/bash
# usage: tools/gpu_prof.sh <tag> <kernel regex> [launch count] — parity tests, breakdown, short bench, ncu --set full (with source) of the named kernel in one 8M wave
tag=${1:-x}; kern=${2:-k_path}; cnt=${3:-2}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
BF_WAVE_LANES=8388608 BF_WAVE_STREAMS=1 timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
    -k regex:$kern --launch-skip 1 --launch-count $cnt python tools/kernel_breakdown.py 8388608 > gpurun_out/ncu_full_$tag.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$tag.log | tail -30
cat gpurun_out/breakdown_8M_$tag.log
python -c "
import json; d = json.load(open('gpurun_out/bench_$tag.json')); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'], d['clocks'])"
tail -3 gpurun_out/bench_$tag.err
