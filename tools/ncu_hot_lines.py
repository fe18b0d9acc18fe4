#!/usr/bin/env python
"""Top source lines of one kernel in an ncu report by warp instructions executed and stall samples.
  python tools/ncu_hot_lines.py report.ncu-rep k_shade [launch_skip] [top_n]"""
import collections, csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--launch-skip", skip,
                      "--launch-count", "1", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fname, hdr = None, None
agg = collections.defaultdict(lambda: [0, 0, 0, ""])   # inst, samples, thread inst, text
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Name":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); ti = hdr.index("Thread Instructions Executed"); continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0]:
        cur = (fname, int(r[0])); agg[cur][3] = r[1].strip()
        continue
    try:
        a = agg[cur]; a[0] += int(r[ii] or 0); a[1] += int(r[si] or 0); a[2] += int(r[ti] or 0)
    except (ValueError, NameError):
        pass
tot_i = sum(a[0] for a in agg.values()); tot_s = sum(a[1] for a in agg.values())
print(f"# {kern} (launch skip {skip}): {tot_i} warp instructions, {tot_s} stall samples")
byfile = collections.Counter()
for (f, l), a in agg.items():
    byfile[f] += a[0]
print("# by file:", ", ".join(f"{f} {v / max(tot_i, 1):.2f}" for f, v in byfile.most_common(8)))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{a[0] / max(tot_i, 1):6.3f} inst {a[1] / max(tot_s, 1):6.3f} stall  lanes {a[2] / max(a[0], 1):5.1f}  {f}:{l}  {a[3][:110]}")
