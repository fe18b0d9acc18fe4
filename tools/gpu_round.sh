#!/bin/bash
# usage: tools/gpu_round.sh <tag> [full]  — one GPU box visit: parity tests, bench (both arms), ncu launch list, ncu --set full of one wave
tag=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
python tools/kernel_breakdown.py 8388608 > gpurun_out/breakdown_8M_$tag.log 2>&1
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --samples 16777216 --no-cpu-baseline > gpurun_out/ncu_launches_$tag.log 2>&1
if [ "$2" = "full" ]; then
  BF_WAVE_LANES=8388608 BF_WAVE_STREAMS=1 timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/prof_$tag \
      --launch-skip 20 --launch-count 20 python tools/kernel_breakdown.py 8388608 > gpurun_out/ncu_full_$tag.log 2>&1
fi
tail -3 gpurun_out/pytest_gpu_$tag.log
cat gpurun_out/breakdown_8M_$tag.log
cat gpurun_out/bench_$tag.json
