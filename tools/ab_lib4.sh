#!/bin/bash
# A/B of alternative library builds on configs c2 / c3 / c5: tools/ab_lib4.sh _alt/libX.so ...
for v in "" "$@"; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  echo "== ${v:-default}"
  for c in c2 c3 c5; do
    python bench.py --config $c --steps 3 --warmup 2 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$c value %.4e' % d['value'], d['roofline']['kernel_ms_per_step'])"
  done
done
