#!/bin/bash
# A/B of alternative library builds on c2 (8 steps) and c3: tools/ab_lib7.sh _alt/libX.so ...
for v in "" "$@" ""; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  for c in c2 c3; do
    python bench.py --config $c --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('${v:-default} $c value %.4e ms %.3f primary %.3f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step']['primary']))"
  done
done
