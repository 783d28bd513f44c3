#!/usr/bin/env python
"""Hot source lines of one launch in an .ncu-rep: stall samples aggregated per CUDA source line.

    python tools/ncu_lines.py rep.ncu-rep <launch index> [top N]
"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, idx = sys.argv[1], int(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass",
                          "--csv", "--launch-skip", str(idx), "--launch-count", "1"],
                         check=True, capture_output=True, text=True).stdout
    agg = collections.Counter()
    inst, src = {}, {}
    f, h, wi, ii = None, None, 0, 0
    for r in csv.reader(io.StringIO(txt)):
        if r and r[0] == "File Name":
            f = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            h = r
            wi = h.index("Warp Stall Sampling (All Samples)")
            ii = h.index("Instructions Executed")
            continue
        if h and len(r) > wi and r[0] != "":
            try:
                w = float(r[wi] or 0)
            except ValueError:
                continue
            key = (f, int(r[0]))
            agg[key] += w
            src[key] = r[1]
            inst[key] = r[ii]
    tot = sum(agg.values())
    print("total samples", tot)
    for k, w in agg.most_common(top):
        print("%5.1f%% %s:%d %s | inst %s" % (100 * w / tot, k[0], k[1], src[k][:80].strip(), inst[k]))


if __name__ == "__main__":
    main()
