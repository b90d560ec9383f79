#!/usr/bin/env python
"""Per-instruction warp-stall table of the hottest loop of an ncu capture taken with --import-source on:
    ncu -i X.ncu-rep --page source --csv > X.csv ; python tools/ncu_loop_table.py X.csv [--all] > table.txt
The loop is found as the address range of the backward branch with the most executions; one line per SASS
instruction: index, samples, cumulative share, instruction, top two stall reasons."""
import csv
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
    stall = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    addr = [int(r[col["Address"]], 16) for r in body]
    smp = [int(r[col["# Samples"]] or 0) for r in body]
    exe = [int(r[col["Instructions Executed"]] or 0) for r in body]
    lo, hi = 0, len(body) - 1
    if "--all" not in sys.argv:
        best = None
        for i, r in enumerate(body):
            m = re.search(r"\bBRA\s+(?:\S+,\s*)?(0x[0-9a-f]+)", r[col["Source"]])
            if m:
                tgt = int(m.group(1), 16)
                if tgt < addr[i] and (best is None or exe[i] > best[0]) and addr[i] - tgt > 0x800:
                    best = (exe[i], tgt, i)
        lo = addr.index(best[1]); hi = best[2]
    tot = sum(smp[lo:hi + 1])
    print("# kernel samples %d, loop samples %d (%.1f %%), loop instructions %d, executions of the back edge %d" %
          (sum(smp), tot, 100.0 * tot / max(1, sum(smp)), hi - lo + 1, exe[hi]))
    cum = 0
    for i in range(lo, hi + 1):
        r = body[i]
        cum += smp[i]
        top = sorted(((int(r[col[n]] or 0), n[6:]) for n in stall), reverse=True)[:2]
        print("%4d %6d %5.3f %-70s %s" % (i - lo, smp[i], cum / max(1, tot), r[col["Source"]][:70],
                                          " ".join("%s:%d" % (n, c) for c, n in top if c)))


if __name__ == "__main__":
    main()
