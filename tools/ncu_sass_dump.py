#!/usr/bin/env python
"""Per-SASS-instruction listing of one kernel launch in an ncu report: index, executed share, active lanes, samples, text.
  python tools/ncu_sass_dump.py report.ncu-rep k_trace [block] [from] [to]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
lo = int(sys.argv[4]) if len(sys.argv) > 4 else 0
hi = int(sys.argv[5]) if len(sys.argv) > 5 else 1 << 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--print-source", "sass"],
                     capture_output=True, text=True).stdout
blocks, cur = [], None
for r in csv.reader(out.splitlines()):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if r and r[0] == "Address":
        cur["hdr"] = r; continue
    if cur is not None and "hdr" in cur and len(r) >= len(cur["hdr"]):
        cur["rows"].append(r)
b = blocks[which]; hdr = b["hdr"]
ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); ti = hdr.index("Thread Instructions Executed")
tot_i = sum(int(r[ii] or 0) for r in b["rows"]); tot_s = sum(int(r[si] or 0) for r in b["rows"])
print(f"# {b['name'][:90]}: {tot_i} warp inst, {tot_s} samples")
for idx, r in enumerate(b["rows"]):
    if lo <= idx < hi:
        n = int(r[ii] or 0)
        print(f"{idx:5d} {n / max(tot_i, 1) * 100:6.3f}% lanes {int(r[ti] or 0) / max(n, 1):5.1f} samp {int(r[si] or 0):5d}  {r[1][:100]}")
