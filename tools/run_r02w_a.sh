#!/bin/bash
# GPU run A of r02w: parity suite with the new anchors + L2 persistence A/B on the two large scenes
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02w_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02w_pytest_gpu.txt
tail -5 gpurun_out/r02w_pytest_gpu.txt
for mb in 0 24 64; do
  BF_L2_PERSIST_MB=$mb timeout 300 python bench.py --config c5 --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02w_c5_l2_$mb.json 2> gpurun_out/r02w_c5_l2_$mb.err
  BF_L2_PERSIST_MB=$mb timeout 300 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02w_c3_l2_$mb.json 2> gpurun_out/r02w_c3_l2_$mb.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r02w_c*_l2_*.json")):
    for line in open(f):
        if line.startswith("{"):
            d=json.loads(line); print(f, "%.4g"%d["value"], d["ms_per_step"], d["roofline"].get("kernel_ms_per_step"))
PY
