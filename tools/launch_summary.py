"""Group an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
Usage: python tools/launch_summary.py launches.csv > summary.csv"""
import csv
import sys
from collections import defaultdict

rows = [ln for ln in open(sys.argv[1]) if ln.startswith('"')]
rd = csv.DictReader(rows)
acc = defaultdict(list)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    name = r["Kernel Name"].split("(")[0]
    acc[name].append(us)
total = sum(sum(v) for v in acc.values())
print("kernel,launches,mean_us,min_us,share_of_launch_list_pct")
for k, v in sorted(acc.items(), key=lambda kv: -sum(kv[1])):
    print(f'"{k}",{len(v)},{sum(v) / len(v):.2f},{min(v):.2f},{100 * sum(v) / total:.1f}')
