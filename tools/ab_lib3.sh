#!/bin/bash
# A/B of alternative library builds on config c3 (mesh-heavy: traversal dominates) and c5
for v in "" "$@"; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  echo "== ${v:-default}"
  for c in c3 c5; do
    python bench.py --config $c --steps 3 --warmup 2 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$c value %.4e' % d['value'], d['roofline']['kernel_ms_per_step'])"
  done
done
