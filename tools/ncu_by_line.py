#!/usr/bin/env python
"""Warp-stall samples per CUDA source line of an ncu capture taken with --import-source on:
    ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > X.csv ; python tools/ncu_by_line.py X.csv [top]
Prints the lines with the most samples and the share of each source file."""
import csv
import sys
from collections import defaultdict


def main():
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    fname, col = None, None
    lines, per_file = [], defaultdict(int)
    for r in csv.reader(open(sys.argv[1])):
        if not r:
            continue
        if r[0] == "File Name":
            fname = r[1].split("/")[-1]
        elif r[0] == "Line No":
            col = {n: i for i, n in enumerate(r)}
        elif col and r[0].isdigit() and len(r) > col["# Samples"]:
            try:
                n = int(r[col["# Samples"]])
            except ValueError:
                continue
            if n:
                stall = sorted(((int(r[i] or 0), k[6:]) for k, i in col.items()
                                if k.startswith("stall_") and "Not Issued" not in k and r[i] not in ("", "-")), reverse=True)[:2]
                lines.append((n, fname, int(r[0]), r[1].strip()[:90], " ".join("%s:%d" % (k, c) for c, k in stall if c)))
                per_file[fname] += n
    tot = sum(per_file.values())
    print("# samples attributed to source lines: %d" % tot)
    for f, n in sorted(per_file.items(), key=lambda x: -x[1]):
        print("# %-24s %7d %5.1f %%" % (f, n, 100.0 * n / tot))
    for n, f, ln, src, st in sorted(lines, reverse=True)[:top]:
        print("%6d %5.2f%% %s:%d  %s   [%s]" % (n, 100.0 * n / tot, f, ln, src, st))


if __name__ == "__main__":
    main()
