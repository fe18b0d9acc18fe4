#!/bin/bash
# r02z: k_path parking-area size chosen per wave slot (path_stash_cap): parity suite, then BF_PATH_STASH_MAX A/B on C2 / C3
python -m pytest tests -m gpu -x -q > gpurun_out/r02z_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02z_pytest_gpu.txt; tail -3 gpurun_out/r02z_pytest_gpu.txt
for m in 8192 65536 131072 262144; do
  for c in c2 c3; do
    BF_PATH_STASH_MAX=$m python bench.py --config $c --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('stash_max $m $c value %.4e ms %.3f primary %.3f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step']['primary']))"
  done
done
