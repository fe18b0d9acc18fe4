#!/bin/bash
# A/B of alternative library builds on config c4 (one k_mega_batch launch per step): tools/ab_lib_c4.sh _alt/libX.so ...
for v in "" "$@"; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  echo "== ${v:-default}"
  python bench.py --config c4 --steps 50 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('c4 value %.4e e2e %.4e ms %.4f' % (d['value'], d['e2e']['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step']['primary'], d['roofline']['kernel'][:40])"
done
