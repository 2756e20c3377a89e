"""Summarise an ncu source page: samples at mbarrier waits by barrier, UTCHMMA, LDTM... (development aid)
usage: python scripts/ncu_waits.py report.ncu-rep"""
import csv, subprocess, sys, io, re
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, data = rows[1], rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
S = [int(r[ci['# Samples']]) for r in data]
src = [r[ci['Source']].strip() for r in data]
tot = sum(S)
print("total samples", tot)
# attribute the BRA after a TRYWAIT to that wait
waits = {}
for i, s in enumerate(src):
    m = re.search(r"TRYWAIT.*\+0x([0-9a-f]+)\]", s)
    if m:
        off = int(m.group(1), 16)
        n = S[i] + (S[i + 1] if i + 1 < len(S) else 0) + (S[i - 1] if 'YIELD' in src[i - 1] else 0)
        waits[off] = waits.get(off, 0) + n
for off in sorted(waits):
    print(f"  wait @0x{off:x}: {waits[off]:7d} samples ({100.0*waits[off]/tot:.1f}%)")
for key in ("UTCHMMA", "LDTM", "UTMALDG", "UTCBAR", "FMNMX"):
    n = sum(S[i] for i, s in enumerate(src) if key in s)
    print(f"  {key}: {n} ({100.0*n/tot:.1f}%)")
