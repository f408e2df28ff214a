#!/usr/bin/env python3
"""Aggregate an ncu source-page (SASS) CSV per CUDA source line.

  ncu -i rep.ncu-rep --page source --csv --kernel-name regex:<k> > src.csv
  cuobjdump -xelf all lib.so ; nvdisasm -g -c <file>.cubin > k.sass
  python tools/ncu_lines.py src.csv k.sass <mangled-function-substring> [top]

ncu's SASS page carries samples / executed-instruction counts per instruction address; nvdisasm -g
carries the file:line of every address.  Joined, they give the share of warp-stall samples and
of executed instructions per source line (profiles/*.lines.txt are made with this).
"""
import csv
import re
import sys
from collections import defaultdict


def line_map(sass_path, func):
    m, cur, active = {}, None, False
    for ln in open(sass_path, errors="replace"):
        if ln.startswith(".text."):
            active = func in ln
            continue
        if not active:
            continue
        f = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if f:
            cur = (f.group(1).split("/")[-1], int(f.group(2)))
            continue
        a = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(\S+)", ln)
        if a:
            m[int(a.group(1), 16)] = (cur, a.group(2))
    return m


def main():
    src_csv, sass, func = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lm = line_map(sass, func)
    rows = list(csv.reader(open(src_csv)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ia, isamp, iex, isrc = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
    stall_cols = [(i, c) for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
    per = defaultdict(lambda: [0, 0, 0, defaultdict(int)])  # samples, inst, fp64 inst, stalls
    base = None
    tot_s = tot_i = 0
    for r in rows[h + 1:]:
        if len(r) <= iex or not r[ia]:
            continue
        try:
            addr = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
        except ValueError:
            continue
        if base is None:
            base = addr
        key, _ = lm.get(addr - base, (None, None))
        s = int(float(r[isamp] or 0)); n = int(float(r[iex] or 0))
        e = per[key]
        e[0] += s; e[1] += n
        if re.match(r"\s*(@!?U?P\d+\s+)?D(FMA|MUL|ADD|SETP|MNMX)|\s*(@!?U?P\d+\s+)?MUFU\.(RSQ|RCP)64H", r[isrc]):
            e[2] += n
        for i, c in stall_cols:
            v = int(float(r[i] or 0))
            if v:
                e[3][c] += v
        tot_s += s; tot_i += n
    print("total samples %d, warp instructions %d" % (tot_s, tot_i))
    print("%-28s %8s %6s %12s %6s %12s  top stalls" % ("file:line", "samples", "%", "warp inst", "%", "fp64 inst"))
    for key, e in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        st = ", ".join("%s %d" % (c[6:], v) for c, v in sorted(e[3].items(), key=lambda kv: -kv[1])[:3])
        name = "%s:%d" % key if key else "?"
        print("%-28s %8d %6.2f %12d %6.2f %12d  %s" % (name, e[0], 100.0 * e[0] / max(tot_s, 1), e[1], 100.0 * e[1] / max(tot_i, 1), e[2], st))


if __name__ == "__main__":
    main()
