#!/bin/bash
# usage: tools/gpu_round.sh <tag> — one GPU box visit for the committed record of a build: parity tests, bench (both arms), ncu launch list
# of the bench command, ncu --set full (with source) of every kernel of one 8M-sample wave (17 launches: k_path, 4 x k_trace closest, 3 x k_trace any-hit, 8 x k_shade, k_finalize)
tag=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --samples 67108864 --no-cpu-baseline --no-extras > gpurun_out/ncu_launches_$tag.log 2>&1
BF_WAVE_LANES=8388608 BF_WAVE_STREAMS=1 timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
    --launch-skip 17 --launch-count 17 python tools/kernel_breakdown.py 8388608 > gpurun_out/ncu_full_$tag.log 2>&1
grep -E "passed|failed" gpurun_out/pytest_gpu_$tag.log | tail -2
cat gpurun_out/breakdown_8M_$tag.log
python -c "
import json; d = json.load(open('gpurun_out/bench_$tag.json')); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'], d['clocks']); print(d.get('small_calls')); print({k: v.get('value') for k, v in (d.get('cpu_rows') or {}).items()})"
