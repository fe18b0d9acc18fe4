#!/bin/bash
# A/B of environment settings on configs c2 / c3 / c5: tools/ab_env4.sh "BF_X=1" ...
for v in "" "$@"; do
  echo "== ${v:-default}"
  for c in c2 c3 c5; do
    env $v python bench.py --config $c --steps 3 --warmup 2 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$c value %.4e' % d['value'], d['roofline']['kernel_ms_per_step'])"
  done
done
