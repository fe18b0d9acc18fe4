#!/usr/bin/env python
"""Top SASS instructions of a kernel in an ncu report by stall samples, with the opcode mix.
  python tools/ncu_sass_hot.py report.ncu-rep k_trace [top_n]"""
import collections, csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# several launches may follow each other: keep the first block
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if r and r[0] == "Address":
        cur["hdr"] = r; continue
    if cur is not None and "hdr" in cur and len(r) >= len(cur["hdr"]):
        cur["rows"].append(r)
which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
b = blocks[which]
hdr = b["hdr"]; ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); ti = hdr.index("Thread Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot_i = sum(int(r[ii] or 0) for r in b["rows"]); tot_s = sum(int(r[si] or 0) for r in b["rows"])
print(f"# {b['name'][:80]} block {which}/{len(blocks)}: {len(b['rows'])} SASS instructions, {tot_i} warp instructions executed, {tot_s} samples")
ops = collections.Counter(); ops_s = collections.Counter()
for r in b["rows"]:
    t = r[1].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    ops[op] += int(r[ii] or 0); ops_s[op] += int(r[si] or 0)
print("# opcode mix (inst share / sample share): " + ", ".join(f"{k} {v / max(tot_i, 1):.3f}/{ops_s[k] / max(tot_s, 1):.3f}" for k, v in ops.most_common(14)))
reasons = collections.Counter()
for r in b["rows"]:
    for c in stall_cols:
        reasons[hdr[c]] += int(r[c] or 0)
print("# stall reasons: " + ", ".join(f"{k[6:]} {v / max(tot_s, 1):.3f}" for k, v in reasons.most_common(8)))
for idx, r in sorted(enumerate(b["rows"]), key=lambda kv: -int(kv[1][si] or 0))[:top]:
    why = max(stall_cols, key=lambda c: int(r[c] or 0))
    print(f"{int(r[si]) / max(tot_s, 1):6.3f} samp {int(r[ii] or 0) / max(tot_i, 1):6.4f} inst  lanes {int(r[ti] or 0) / max(int(r[ii] or 0), 1):5.1f}  #{idx:5d} {hdr[why][6:]:14s} {r[1][:90]}")
