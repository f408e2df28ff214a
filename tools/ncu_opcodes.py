#!/usr/bin/env python3
"""Executed-instruction histogram by SASS opcode from an ncu source-page CSV.

  ncu -i rep.ncu-rep --page source --csv --kernel-name regex:<k> > src.csv
  python tools/ncu_opcodes.py src.csv [top]
"""
import collections
import csv
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    for i, r in enumerate(rows):
        if "Source" in r and "Instructions Executed" in r:
            hdr, start = r, i + 1
            break
    si, ei = hdr.index("Source"), hdr.index("Instructions Executed")
    cnt, full, tot = collections.Counter(), collections.Counter(), 0
    for r in rows[start:]:
        try:
            n = int(r[ei])
        except (ValueError, IndexError):
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[si])
        if not m:
            continue
        cnt[m.group(2).split(".")[0]] += n
        full[m.group(2)] += n
        tot += n
    print("warp instructions executed:", tot)
    for op, n in cnt.most_common(top):
        print("%-10s %12d %5.1f%%" % (op, n, 100.0 * n / tot))
    print("-- with modifiers")
    for op, n in full.most_common(top):
        print("%-24s %12d %5.1f%%" % (op, n, 100.0 * n / tot))


if __name__ == "__main__":
    main()
