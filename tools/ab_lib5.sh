#!/bin/bash
# A/B of alternative library builds on c2 / c3 (5 steps): tools/ab_lib5.sh _alt/libX.so ...
for v in "" "$@" ""; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  for c in c2 c3; do
    python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('${v:-default} $c value %.4e ms %.3f' % (d['value'], d['ms_per_step']))"
  done
done
