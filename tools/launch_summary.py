#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): per-kernel count, mean us, share."""
import csv, re, sys
from collections import OrderedDict
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = OrderedDict()
for r in rows[1 + skip:]:
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("femb::", "")
    v = float(r[vi].replace(",", ""))
    v = v / 1000.0 if r[ui] in ("ns", "nsecond") else v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print("# %s (first %d launches skipped): %d launches, %.1f us total (cold-cache, serialised: compare SHARES)" % (sys.argv[1], skip, sum(a[0] for a in agg.values()), tot))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-70s n=%4d  mean %9.2f us  share %5.1f%%" % (k[:70], a[0], a[1] / a[0], 100 * a[1] / tot))
