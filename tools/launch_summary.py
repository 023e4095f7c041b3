#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`): python tools/launch_summary.py <csv>"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
h = rows[hdr]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6}
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    name = r[ki].split("(")[0][:64]
    tot[name] += float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    cnt[name] += 1
total = sum(tot.values())
print(f"# {sys.argv[1]}: {sum(cnt.values())} launches, {total / 1e3:.3f} ms of GPU time (cold-cache, serialised: shares, not absolutes)")
for name, v in tot.most_common():
    print(f"{name:66s} {cnt[name]:4d} launches {v:12.1f} us {100 * v / total:6.2f} %")
