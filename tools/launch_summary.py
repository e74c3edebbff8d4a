#!/usr/bin/env python3
"""Per-kernel summary of an ncu launch list (`--metrics gpu__time_duration.sum --csv`): launches, device time, share."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
agg = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if len(r) != len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[ix["Metric Value"]].replace(",", ""))
    u = r[ix["Metric Unit"]]
    us = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
    name = r[ix["Kernel Name"]].split("(")[0]
    agg[name][0] += 1
    agg[name][1] += us
tot = sum(v[1] for v in agg.values())
print("%d launches, %.1f us of device time in total" % (sum(v[0] for v in agg.values()), tot))
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print("%6d launches %10.1f us %5.1f%%  %s" % (n, us, 100 * us / tot, name[:110]))
