"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: python scripts/launch_summary.py file.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hdr, d = None, collections.defaultdict(list)
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        rec = dict(zip(hdr, r))
        if rec.get("Metric Name") == "gpu__time_duration.sum":
            v = float(rec["Metric Value"].replace(",", ""))
            u = rec["Metric Unit"]
            v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
            d[rec["Kernel Name"][:70]].append(v)
tot = sum(sum(v) for v in d.values())
for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
    print("%-72s n=%3d mean=%8.1f us share=%5.1f%%" % (k, len(v), sum(v) / len(v), 100 * sum(v) / tot))
