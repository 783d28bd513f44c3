#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove the tcgen05 / TMA / tensor-memory path
(B200_PROFILING.md): UTCHMMA (tcgen05.mma kind::tf32), UTMALDG (TMA tensor load), LDTM / STTM
(tcgen05.ld / st), UTCBAR (tcgen05.commit), SYNCS (mbarrier try_wait / arrive), plus FFMA for the
exact kernels.   usage: python tools/sass_summary.py > profiles/sass_summary.txt"""
import collections
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "hnsw_ivf_b200/lib/libb200ivf.so"
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
COLS = ("UTCHMMA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "SYNCS", "FFMA")
ins = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)")
rows, name, cnt = [], None, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if name:
            rows.append((name, cnt))
        name, cnt = m.group(1), collections.Counter()
        continue
    m = ins.match(line)
    if m and name:
        cnt[m.group(1)] += 1
        cnt["total"] += 1
if name:
    rows.append((name, cnt))
names = subprocess.run(["c++filt"], input="\n".join(n for n, _ in rows), capture_output=True,
                       text=True).stdout.splitlines()
print("# cuobjdump -sass %s (sm_100a): instruction counts per kernel" % so)
print("# %-86s %s total" % ("kernel", " ".join(COLS)))
for (n, c), dn in sorted(zip(rows, names), key=lambda t: t[1]):
    dn = re.sub(r"\(anonymous namespace\)::|b200ivf::|^void ", "", dn)
    dn = re.sub(r"\(.*$", "", dn)
    print("%-88s %s %d" % (dn[:88], " ".join("%7d" % c[k] for k in COLS), c["total"]))
