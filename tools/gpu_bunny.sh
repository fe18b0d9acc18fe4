#!/bin/bash
# usage: tools/gpu_bunny.sh <tag> — C2 with the stand-in SURVEY 8(d) names (resources/data/common/meshes/bunny.ply, copied to _alt/ where the
# reference tree exists) beside the procedural hull: bench line + per-kernel breakdown of one step
tag=${1:-x}
mkdir -p gpurun_out
python bench.py --mesh _alt/bunny.ply --steps 10 --no-cpu-baseline --no-extras > gpurun_out/bench_c2_bunny_$tag.json 2> gpurun_out/bench_c2_bunny_$tag.err; tail -2 gpurun_out/bench_c2_bunny_$tag.err
python bench.py --steps 10 --no-cpu-baseline --no-extras > gpurun_out/bench_c2_hull_$tag.json 2> gpurun_out/bench_c2_hull_$tag.err
python - <<PY
import json
for n in ("bunny", "hull"):
    d = json.load(open(f"gpurun_out/bench_c2_{n}_$tag.json"))
    print(n, d["config"]["triangles"], "value %.4e e2e %.4e" % (d["value"], d["e2e"]["value"]), d["roofline"]["kernel_ms_per_step"], d["roofline"]["queries_per_sample"], "build_ms", d["run"]["build_ms"])
PY
