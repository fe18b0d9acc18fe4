#!/bin/bash
# one-GPU bench lines of the other configs for the final build (tag r02x)
for c in c1 c3 c4 c5; do
  python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/bench_${c}_r02x.json 2> gpurun_out/bench_${c}_r02x.err; tail -1 gpurun_out/bench_${c}_r02x.err
done
python - <<PY
import json
for c in ("c1","c3","c4","c5"):
    try:
        d=json.load(open(f"gpurun_out/bench_{c}_r02x.json"))
        print(c, "value %.3e e2e %.3e ms/step %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "items/s", d.get("items_per_s"), "launches", d["gpu_launches"], "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(c, "FAILED", e)
PY
