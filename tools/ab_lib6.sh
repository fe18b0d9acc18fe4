#!/bin/bash
# A/B of alternative library builds on c3 / c5 (and c2): tools/ab_lib6.sh _alt/libX.so ...
for v in "" "$@"; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  for c in c3 c5 c2; do
    st=5; [ $c = c5 ] && st=3
    python bench.py --config $c --steps $st --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('${v:-default} $c value %.4e ms %.3f shade %.3f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step']['shade']))"
  done
done
