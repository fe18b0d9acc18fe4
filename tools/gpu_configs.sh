#!/bin/bash
# usage: tools/gpu_configs.sh <tag> — parity tests, then the bench line of every config (short runs), both arms for c2
tag=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$tag.log | tail -30
python bench.py > gpurun_out/bench_c2_$tag.json 2> gpurun_out/bench_c2_$tag.err; tail -2 gpurun_out/bench_c2_$tag.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c2_reference_$tag.json 2> gpurun_out/bench_c2_reference_$tag.err
for c in c1 c3 c4 c5; do
  python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/bench_${c}_$tag.json 2> gpurun_out/bench_${c}_$tag.err; tail -2 gpurun_out/bench_${c}_$tag.err
done
python - <<PY
import json
for c in ("c2","c1","c3","c4","c5"):
    try:
        d=json.load(open(f"gpurun_out/bench_{c}_$tag.json"))
        print(c, "value %.3e e2e %.3e ms/step %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "items/s", d.get("items_per_s"), "launches", d["gpu_launches"], "cpu", (d.get("cpu_baseline") or {}).get("value"))
        print("   roofline", {k: d["roofline"][k] for k in ("frac","kernel_share_of_step","avg_launch_ms")}, d["roofline"]["kernel_ms_per_step"])
        if "small_calls" in d: print("   small", d["small_calls"])
        if "cpu_rows" in d: print("   cpu_rows", {k:v.get("value") for k,v in d["cpu_rows"].items()})
    except Exception as e:
        print(c, "FAILED", e)
print(open("gpurun_out/bench_c2_reference_$tag.json").read()[:300])
PY
