#!/bin/bash
# usage: tools/gpu_quick.sh <tag> [pytest args]  — parity tests, per-kernel breakdown of one 8M-lane wave, short bench (no CPU arm)
tag=${1:-x}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -q "$@" > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$tag.log | tail -30
cat gpurun_out/breakdown_8M_$tag.log
python -c "
import json; d = json.load(open('gpurun_out/bench_$tag.json')); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'], d['clocks'])"
tail -3 gpurun_out/bench_$tag.err
