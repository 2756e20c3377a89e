"""Warp-stall reasons of one kernel, summed over its SASS, from an `ncu --page source --csv` dump (development aid).
usage: python scripts/ncu_stall_reasons.py source.csv [line_lo line_hi]   (optional CUDA source line window)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], newline="")))
hdr = next(i for i, r in enumerate(rows) if any(c.startswith("stall_") for c in r))
h = rows[hdr]
cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
lo, hi = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (None, None)
tot = {c: 0.0 for _, c in cols}
cur = None
for r in rows[hdr + 1:]:
    if len(r) < len(h):
        continue
    if not r[0].strip().isdigit():
        continue  # SASS rows repeat what their CUDA source line row already aggregates
    cur = int(r[0])
    if lo is not None and not (lo <= cur <= hi):
        continue
    for i, c in cols:
        try:
            tot[c] += float(r[i].replace(",", "") or 0)
        except ValueError:
            pass
s = sum(tot.values()) or 1.0
print(f"samples {s:.0f}" + (f" in source lines {lo}-{hi}" if lo is not None else ""))
for c, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print(f"  {c:28s} {v:10.0f} {100 * v / s:6.2f}%")
