#!/usr/bin/env python
"""Print a per-launch table from a tools/gpu_launchlist.sh csv."""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith("==")]
h = rows[0]
kn, mn, mv, idc = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
d = {}
for r in rows[1:]:
    if len(r) > mv:
        d.setdefault((int(r[idc]), r[kn][:28]), {})[r[mn].split("__")[-1][:28]] = r[mv]
for k in sorted(d):
    print(k[0], k[1], " ".join(f"{a}={b}" for a, b in d[k].items()))
