#!/bin/bash
# fixed cost per bf_receive call on C2: step time against request size
for lg in 26 27 28 29 30; do
  n=$((1<<lg))
  timeout 300 python bench.py --config c2 --samples $n --steps 6 --warmup 3 --no-cpu-baseline --no-extras 2> gpurun_out/r02w_b_$lg.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print($lg, d['ms_per_step'], '%.4g'%d['value'], d['roofline'].get('kernel_ms_per_step'))
"
done
