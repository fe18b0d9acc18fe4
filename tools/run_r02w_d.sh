#!/bin/bash
# tapered wave schedule A/B on C2 (and C3)
run() { # env..., config, samples
  env "$1" "$2" timeout 300 python bench.py --config $3 --samples $4 --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2> gpurun_out/r02w_d.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1 $2 $3 $4', d['ms_per_step'], '%.4g'%d['value'])
"; }
for rep in 1 2; do
run BF_WAVE_TAPER=0 X=0 c2 268435456
run BF_WAVE_TAPER=1 X=0 c2 268435456
done
run BF_WAVE_TAPER=1 BF_WAVE_LANES=16777216 c2 268435456
run BF_WAVE_TAPER=0 BF_WAVE_LANES=16777216 c2 268435456
run BF_WAVE_TAPER=0 X=0 c2 536870912
run BF_WAVE_TAPER=1 X=0 c2 536870912
run BF_WAVE_TAPER=0 X=0 c3 0
run BF_WAVE_TAPER=1 X=0 c3 0
