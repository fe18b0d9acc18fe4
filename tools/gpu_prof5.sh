#!/bin/bash
# usage: tools/gpu_prof5.sh <tag> — C5 (1e7 triangles): per-kernel times of one 2^20-sample wave, then ncu (selected metrics) of the same wave
tag=$1
mkdir -p gpurun_out
python tools/c5_breakdown.py 1048576 > gpurun_out/c5_breakdown_$tag.txt 2>&1; cat gpurun_out/c5_breakdown_$tag.txt
timeout 1200 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum \
  --clock-control none -c 60 --csv --log-file gpurun_out/c5_metrics_$tag.csv python tools/c5_breakdown.py 1048576 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/c5_metrics_$tag.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
out={}
for r in rows[1:]:
    out.setdefault((int(r[ii]), r[ki].split("(")[0][-26:]), {})[r[mi]]=r[vi]
short={"gpu__time_duration.sum":"ns","smsp__inst_executed.sum":"inst","smsp__thread_inst_executed_per_inst_executed.ratio":"thr/inst","smsp__issue_active.avg.pct_of_peak_sustained_active":"issue%","sm__warps_active.avg.pct_of_peak_sustained_active":"warps%","lts__t_sector_hit_rate.pct":"L2hit","l1tex__t_sector_hit_rate.pct":"L1hit","dram__bytes_read.sum":"dramR","dram__bytes_write.sum":"dramW","smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio":"lsb","l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum":"ldsect","l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum":"ldreq"}
with open("gpurun_out/c5_metrics_$tag.txt","w") as f:
    for k,v in sorted(out.items()):
        line=f"{k[0]:3d} {k[1]:26s} " + " ".join(f"{short.get(a,a)}={b}" for a,b in v.items())
        print(line); f.write(line+"\n")
PY
