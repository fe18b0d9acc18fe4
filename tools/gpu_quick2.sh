#!/bin/bash
# usage: tools/gpu_quick2.sh <tag> — parity tests, C4 / C1 bench lines, small-call sweep
tag=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$tag.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$tag.log | tail -30
for c in c4 c5; do
  python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${c}_$tag.json 2> gpurun_out/bench_${c}_$tag.err; tail -2 gpurun_out/bench_${c}_$tag.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_${c}_$tag.json')); print('$c', 'value %.3e e2e %.3e ms/step %.2f' % (d['value'], d['e2e']['value'], d['ms_per_step']), 'items/s', d.get('items_per_s'), d['e2e'].get('items_per_s'), 'launches', d['gpu_launches'])"
done
python tools/small_calls.py 2>&1 | tee gpurun_out/small_calls_$tag.txt
