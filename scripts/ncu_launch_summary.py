#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (share of the summed time)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[i]
ix = {k: j for j, k in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[i + 1:]:
    if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    key = (r[ix["Kernel Name"]].split("(")[0][:60], r[ix["Grid Size"]], r[ix["Block Size"]])
    agg.setdefault(key, []).append(float(r[ix["Metric Value"]]) / 1e3)
total = sum(sum(v) for v in agg.values())
print(f"{'kernel':60s} {'grid':>16s} {'block':>14s} {'n':>4s} {'mean us':>10s} {'min us':>10s} {'share':>7s}")
for (k, g, b), v in agg.items():
    print(f"{k:60s} {g:>16s} {b:>14s} {len(v):4d} {sum(v)/len(v):10.1f} {min(v):10.1f} {100*sum(v)/total:6.1f}%")
print(f"total {total/1e3:.3f} ms over {sum(len(v) for v in agg.values())} launches")
