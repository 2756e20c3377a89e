"""Hottest CUDA source lines (and SASS instructions) of an `ncu --page source --print-source cuda,sass --csv` dump,
ranked by sampled warp stalls.  usage: python scripts/ncu_hot_lines.py source.csv [n]"""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1], newline="")))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = next((i for i, r in enumerate(rows) if any("Sampling" in c for c in r)), None)
if hdr is None:
    print("no sampling columns; first rows:", rows[:3]); sys.exit(0)
h = rows[hdr]
print("columns:", h)
si = next(i for i, c in enumerate(h) if "Sampling" in c and "All" in c)
ii = next((i for i, c in enumerate(h) if c.strip() == "Instructions Executed"), None)
srci = next((i for i, c in enumerate(h) if c.strip() == "Source"), 1)
body = []
cur_line = ""
by_line = defaultdict(lambda: [0.0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= si:
        continue
    try:
        v = float(r[si].replace(",", "") or 0)
    except ValueError:
        continue
    ins = 0.0
    if ii is not None:
        try: ins = float(r[ii].replace(",", "") or 0)
        except ValueError: pass
    body.append((v, ins, r))
tot = sum(b[0] for b in body) or 1.0
toti = sum(b[1] for b in body) or 1.0
print(f"total samples {tot:.0f}, instructions {toti:.0f}")
for v, ins, r in sorted(body, key=lambda b: -b[0])[:n]:
    print(f"{100 * v / tot:6.2f}% stall {100 * ins / toti:6.2f}% inst  {r[0][:18]:18s} {(r[srci].strip() or (r[3].strip() if len(r) > 3 else ""))[:140]}")
