#!/usr/bin/env python
"""profiles/roofline_traffic.json (read by bench.py's roofline block) from the `ncu_summary.py full` text of one C2 wave.
  python tools/roofline_traffic.py profiles/rNN_full_wave8M_all_kernels.txt 8388608"""
import json, os, re, sys
path, samples = sys.argv[1], int(sys.argv[2])
lines = open(path).read().splitlines()
names = [n.strip() for n in re.findall(r'#\d+ [^#]+', lines[1])]
K = len(names)
rows = {}
for l in lines[2:]:
    c = l.split()
    if len(c) >= K + 1:
        rows[c[0]] = c[-K:]
unit = {l.split()[0]: l.split()[1] for l in lines[2:] if len(l.split()) >= K + 2}
def col(m):
    return [float(x) for x in rows[m]]
scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
inst = col("smsp__inst_executed.sum")
dram = [(r * scale[unit["dram__bytes_read.sum"]] + w * scale[unit["dram__bytes_write.sum"]]) for r, w in zip(col("dram__bytes_read.sum"), col("dram__bytes_write.sum"))]
tpi = col("smsp__thread_inst_executed_per_inst_executed.ratio")
ia = col("smsp__issue_active.avg.pct_of_peak_sustained_active")
kp = [i for i, n in enumerate(names) if "k_path" in n]
assert len(kp) == 1, names
kp = kp[0]
trav = [i for i, n in enumerate(names) if "k_trace" in n]
out = {"config": "c2",
       "source": f"{path} (ncu --set full --clock-control none, one {samples:,}-sample wave, {K} launches)".replace(",", " "),
       "k_path_warp_inst_per_sample": inst[kp] / samples, "step_warp_inst_per_sample": sum(inst) / samples,
       "k_path_dram_bytes_per_sample": dram[kp] / samples, "step_dram_bytes_per_sample": sum(dram) / samples,
       "k_path_thread_inst_per_inst": tpi[kp], "k_path_issue_active_pct": ia[kp],
       "traversal_thread_inst_per_inst": sum(tpi[i] * inst[i] for i in trav) / sum(inst[i] for i in trav),
       "launches": [{"kernel": names[i], "warp_inst": inst[i], "thread_inst_per_inst": tpi[i], "issue_active_pct": ia[i], "dram_bytes": dram[i]} for i in range(K)]}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "roofline_traffic.json"), "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k != "launches"}, indent=1))
