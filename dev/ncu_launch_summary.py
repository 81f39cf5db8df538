"""Aggregate an `ncu --csv --metrics gpu__time_duration.sum` launch list per kernel name.
usage: python dev/ncu_launch_summary.py launches.csv [skip_first_n_launches] [steps]"""
import csv
import re
import sys
from collections import OrderedDict

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
lines = open(path, errors="replace").read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[start:]))
rows = [r for r in rows if r.get("Metric Name") == "gpu__time_duration.sum"][skip:]
agg = OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r["Kernel Name"])[:70]
    v = float(r["Metric Value"].replace(",", ""))
    if r["Metric Unit"] in ("ns", "nsecond"):
        v /= 1e3
    elif r["Metric Unit"] in ("ms", "msecond"):
        v *= 1e3
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':72s} {'launches/step':>13s} {'us/step':>9s} {'share':>6s}")
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:72s} {n / steps:13.1f} {us / steps:9.1f} {100 * us / tot:5.1f}%")
print(f"{'total':72s} {len(rows) / steps:13.1f} {tot / steps:9.1f}")
