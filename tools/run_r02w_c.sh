#!/bin/bash
# BVH leaf-size experiment: BF_BVH_SPLIT_SHARE (0.25 = today: nearly every leaf child holds one triangle)
for f in 0.25 0.4 0.6 1.0; do
  for c in c2 c3 c5; do
    st=6; [ $c = c5 ] && st=3
    BF_BVH_SPLIT_SHARE=$f timeout 300 python bench.py --config $c --steps $st --warmup 3 --no-cpu-baseline --no-extras 2> gpurun_out/r02w_c_${c}_$f.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$f $c', d['ms_per_step'], '%.4g'%d['value'], d['run'].get('build_ms'), d['roofline'].get('kernel_ms_per_step'))
"
  done
done
