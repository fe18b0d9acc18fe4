#!/bin/bash
# A/B of alternative builds of the native library: tools/ab_lib.sh _alt/libA.so _alt/libB.so ...  ("" = the default build)
for v in "" "$@"; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  echo "== ${v:-default}"
  python tools/kernel_breakdown.py 8388608 2>&1 | sed -n 3,4p
  python bench.py --steps 10 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'])"
done
