#!/bin/bash
# usage: tools/gpu_multi.sh <tag> <N> — N-GPU box visit: the C++ two-GPU test, then the bench at N (weak, strong, c3 weak, c4, c5)
tag=${1:-x}; N=${2:-2}
mkdir -p gpurun_out
python -m pytest tests/test_cpp_host.py -m gpu -q 2>&1 | tail -3
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
run --steps 10 --warmup 3 > gpurun_out/bench_c2_weak_${N}gpu_$tag.json 2> gpurun_out/bench_c2_weak_${N}gpu_$tag.err
run --steps 10 --warmup 3 --scaling strong > gpurun_out/bench_c2_strong_${N}gpu_$tag.json 2> gpurun_out/bench_c2_strong_${N}gpu_$tag.err
run --config c3 --steps 5 --warmup 3 > gpurun_out/bench_c3_weak_${N}gpu_$tag.json 2> gpurun_out/bench_c3_weak_${N}gpu_$tag.err
run --config c4 --steps 5 --warmup 3 > gpurun_out/bench_c4_${N}gpu_$tag.json 2> gpurun_out/bench_c4_${N}gpu_$tag.err
run --config c5 --steps 3 --warmup 2 > gpurun_out/bench_c5_${N}gpu_$tag.json 2> gpurun_out/bench_c5_${N}gpu_$tag.err
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_*_${N}gpu_$tag.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "value %.4e e2e %.4e ms/step %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), d.get("items_per_s"), d.get("reduce_check"))
    except Exception as e:
        print(f, "FAILED", e); print(open(f.replace(".json", ".err")).read()[-1500:])
PY
