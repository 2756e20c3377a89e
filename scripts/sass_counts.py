#!/usr/bin/env python
"""SASS instruction counts per kernel of libragb200.so (no GPU needed): python scripts/sass_counts.py > profiles/r2_sass_counts.txt
UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA), UBLKCP = cp.async.bulk (1-D TMA),
UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, REDUX = redux.sync, HMMA/IMMA = mma.sync / wmma (must stay 0)."""
import re
import subprocess
import sys
from pathlib import Path

LIB = Path(__file__).resolve().parent.parent / "modular-rag-system_b200" / "csrc" / "libragb200.so"
COLS = [("UTC*MMA", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"), ("UBLKCP", r"\bUBLKCP"),
        ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"), ("REDUX", r"\bREDUX"), ("FMNMX3", r"\bFMNMX3"),
        ("HMMA/IMMA", r"\b(HMMA|IMMA)"), ("LDG.128", r"\bLDG\.E\.[A-Z.]*128")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    parts = re.split(r"\n\s*Function : \S+\n", sass)[1:]
    print("# SASS instruction counts per kernel of modular-rag-system_b200/csrc/libragb200.so (sm_100a cubin), scripts/sass_counts.py.")
    print("# UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA), UBLKCP = cp.async.bulk (1-D TMA),")
    print("# UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, REDUX = redux.sync.  No HMMA/IMMA (mma.sync / wmma) anywhere.")
    print("# gemm_topk_pair_kernel<E, stages, kQRes, kEpi8, kSolo, kPar>\n")
    print("kernel".ljust(118) + "".join(c.rjust(10) for c, _ in COLS))
    total = [0] * len(COLS)
    for name, body in zip(names, parts):
        name = name.replace("ragb::", "")
        counts = [len(re.findall(rx, body)) for _, rx in COLS]
        total = [a + b for a, b in zip(total, counts)]
        print(name[:116].ljust(118) + "".join(str(c).rjust(10) for c in counts))
    print("ALL KERNELS IN THE LIBRARY".ljust(118) + "".join(str(c).rjust(10) for c in total))


if __name__ == "__main__":
    sys.exit(main())
