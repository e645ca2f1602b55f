#!/usr/bin/env python3
"""ncu `--metrics gpu__time_duration.sum --csv` launch list -> per-kernel table (launches, total us, share, us/launch).
    python tools/summarize_launches.py gpurun_out/launches.csv "header line" > profiles/rNN_launches_summary.txt"""
import collections
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1], errors="replace") if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi or r[hdr.index("Metric Name")] != "gpu__time_duration.sum":
        continue
    v = float(r[vi].replace(",", ""))
    us = v / 1e3 if r[ui] in ("ns", "nsecond") else v * 1e3 if r[ui] in ("ms", "msecond") else v
    a = agg.setdefault(r[ki], [0, 0.0])
    a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
for line in sys.argv[2:]:
    print("# " + line)
print(f"{'kernel':60s} {'launches':>8s} {'total_us':>12s} {'share':>7s} {'us/launch':>10s}")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:60]:60s} {n:8d} {us:12.1f} {us / tot * 100:6.2f}% {us / n:10.1f}")
