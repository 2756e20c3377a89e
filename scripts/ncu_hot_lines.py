"""Top source/SASS lines of an `ncu --page source --csv` dump by sampled warp stalls.
usage: python scripts/ncu_hot_lines.py source.csv [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], newline="")))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
for i, r in enumerate(rows):
    if any("Sampling" in c for c in r):
        hdr = i
        break
if hdr is None:
    print("no sampling columns found; header candidates:", rows[:3])
    sys.exit(0)
h = rows[hdr]
si = next(i for i, c in enumerate(h) if "Sampling" in c and "All" in c) if any("Sampling" in c and "All" in c for c in h) else next(i for i, c in enumerate(h) if "Sampling" in c)
src = next((i for i, c in enumerate(h) if c.strip() in ("Source", "SASS", "Instruction")), 1)
body = []
for r in rows[hdr + 1:]:
    try:
        body.append((float(r[si].replace(",", "") or 0), r))
    except (ValueError, IndexError):
        continue
tot = sum(b[0] for b in body) or 1.0
print("columns:", h[:8])
for v, r in sorted(body, key=lambda b: -b[0])[:n]:
    print(f"{100 * v / tot:6.2f}%  {r[0][:12]:12s} {r[src][:150]}")
