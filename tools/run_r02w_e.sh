#!/bin/bash
# packed-SM traversal launches (BF_TRACE_PACK) A/B: parity first, then C2 / C3
BF_TRACE_PACK=1 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c2_full_size or path_kernel or profile_small or sample_records" 2>&1 | tail -3
run() {
  env "$1" "$2" timeout 300 python bench.py --config $3 --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2> gpurun_out/r02w_e.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1 $2 $3', d['ms_per_step'], '%.4g'%d['value'], d['roofline'].get('kernel_ms_per_step'))
"; }
for rep in 1 2; do
run BF_TRACE_PACK=0 X=0 c2
run BF_TRACE_PACK=1 X=0 c2
done
run BF_TRACE_PACK=1 BF_TRACE_ROUNDS=16 c2
run BF_TRACE_PACK=1 BF_TRACE_ROUNDS=96 c2
run BF_TRACE_PACK=0 X=0 c3
run BF_TRACE_PACK=1 X=0 c3
run BF_TRACE_PACK=1 BF_TRACE_ROUNDS=16 c3
tail -5 gpurun_out/r02w_e.err
