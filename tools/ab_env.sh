#!/bin/bash
# A/B of launch-shape environment settings: tools/ab_env.sh "BF_X=1 BF_Y=2" "BF_X=3" ...  ("" = defaults)
for v in "" "$@"; do
  echo "== ${v:-default}"
  env $v python bench.py --steps 6 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'])"
done
