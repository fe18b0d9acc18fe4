#!/usr/bin/env python
"""Global loads of one kernel in an ncu report split by the source line that issues them: L1 tag requests, sectors passed on to L2
(L1 misses), and the share of each line -- the node / triangle split of the traversal's memory traffic `north_star` asks for.
  python tools/ncu_load_split.py report.ncu-rep k_trace [launch_skip]"""
import collections, csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--launch-skip", skip,
                      "--launch-count", "1", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
hdr, cur = None, None
agg = collections.defaultdict(lambda: [0, 0, 0, 0, ""])   # load instructions executed, thread-level, L1 tag requests, L2 sectors, text
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "Line No":
        hdr = r
        ii, ti = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
        l1, l2, op, sp = hdr.index("L1 Tag Requests Global"), hdr.index("L2 Theoretical Sectors Global"), hdr.index("Access Operation"), hdr.index("Address Space")
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0]:
        cur = int(r[0]); agg[cur][4] = r[1].strip()
        continue
    if cur is None or "Global" not in r[sp] or "Load" not in r[op]:
        continue
    a = agg[cur]
    a[0] += int(r[ii] or 0); a[1] += int(r[ti] or 0); a[2] += int(r[l1] or 0); a[3] += int(r[l2] or 0)
rows = [(l, a) for l, a in agg.items() if a[0]]
tot_req = sum(a[2] for _, a in rows); tot_sec = sum(a[3] for _, a in rows)
print(f"# {kern} (launch skip {skip}): global loads by source line; L1 tag requests {tot_req}, sectors requested from L2 {tot_sec}")
print("# line   warp loads   lanes   L1 tag requests (share)   sectors to L2 (share)   sectors per request   source")
for l, a in sorted(rows, key=lambda kv: -kv[1][3])[:14]:
    print(f"{l:5d} {a[0]:12d} {a[1] / a[0]:7.1f} {a[2]:14d} ({a[2] / max(tot_req, 1):5.3f}) {a[3]:14d} ({a[3] / max(tot_sec, 1):5.3f}) {a[3] / max(a[2], 1):8.2f}   {a[4][:100]}")
