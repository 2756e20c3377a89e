"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python scripts/launch_table.py launches.csv [last_n_launches_to_list]"""
import csv, sys
from collections import OrderedDict
rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") == "gpu__time_duration.sum":
        v = float(r["Metric Value"].replace(",", ""))
        u = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
        rows.append((r["Kernel Name"], v))
tot = OrderedDict()
for k, v in rows:
    a = tot.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
s = sum(v for _, v in rows) or 1.0
print(f"{'kernel':95s} {'launches':>8s} {'total_us':>12s} {'share':>7s}")
for k, (n, v) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:95]:95s} {n:8d} {v:12.1f} {100 * v / s:6.1f}%")
n = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if n:
    print(f"-- last {n} launches --")
    for k, v in rows[-n:]:
        print(f"{v:10.1f} us  {k[:110]}")
