#!/bin/bash
# A/B of alternative library builds on the quick GPU parity subset + one big wave + the bench
for v in "" "$@"; do
  if [ -n "$v" ]; then export BF_NATIVE_LIB=/root/repo/$v; else unset BF_NATIVE_LIB; fi
  echo "== ${v:-default}"
  python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -1
  python tools/kernel_breakdown.py 67108864 2>&1 | tail -1
  python bench.py --steps 6 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('value', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['roofline']['kernel_ms_per_step'])"
done
