#!/bin/bash
# usage: tools/gpu_prof3.sh <tag> [lib] — ncu (selected metrics) of the traversal launches of one C3 wave
tag=$1; lib=$2
if [ -n "$lib" ]; then export BF_NATIVE_LIB=/root/repo/$lib; fi
mkdir -p gpurun_out
python tools/c3_wave.py 16777216 > gpurun_out/c3_wave_$tag.txt 2>&1; cat gpurun_out/c3_wave_$tag.txt
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum \
  --clock-control none -k regex:"k_trace|k_primary" --launch-skip 8 --launch-count 8 --csv --log-file gpurun_out/c3_trace_metrics_$tag.csv python tools/c3_wave.py 16777216 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/c3_trace_metrics_$tag.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
out={}
for r in rows[1:]:
    out.setdefault((r[ii], r[ki].split("(")[0][-28:]), {})[r[mi]]=r[vi]
for k,v in out.items():
    print(k, {a.split("__")[1][:28]: b for a,b in v.items()})
PY
