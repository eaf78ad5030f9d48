#!/usr/bin/env python
"""Condensed SASS hot-spot listing from `ncu -i X.ncu-rep --page source --csv`:
   tools/ncu_sass.py rep [kernel_index] [min_pct]"""
import csv, subprocess, sys
rep = sys.argv[1]; kidx = int(sys.argv[2]) if len(sys.argv) > 2 else 0; minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.7
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
kernels, cur = [], None
for row in csv.reader(out.splitlines()):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}; kernels.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = row
    elif cur is not None and len(row) == len(cur["hdr"]):
        cur["rows"].append(row)
k = kernels[kidx]; h = k["hdr"]
si, ii, ti = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
stall = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[si]) for r in k["rows"]); toti = sum(int(r[ii]) for r in k["rows"])
print("#", k["name"][:100]); print("# samples", tot, "warp-instr", toti, "thread-instr", sum(int(r[ti]) for r in k["rows"]))
agg = {}
for r in k["rows"]:
    for i in stall: agg[h[i]] = agg.get(h[i], 0) + int(r[i])
print("# stalls:", ", ".join("%s %.1f%%" % (n[6:], 100 * v / tot) for n, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
cum = 0
for n, r in enumerate(k["rows"]):
    s = int(r[si]); cum += int(r[ii]); op = r[1].strip()
    key = any(t in op for t in ("BAR", "SYNCS", "UBLKCP", "EXIT", "BRA")) 
    if 100.0 * s / tot >= minpct or (key and s > 0):
        top = max(stall, key=lambda i: int(r[i]))
        print("%5d %-58s smp %5.2f%% inst %9s cum %5.1f%% %s" % (n, op[:58], 100.0 * s / tot, r[ii], 100.0 * cum / toti, h[top][6:]))
