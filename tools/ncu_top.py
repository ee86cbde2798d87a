#!/usr/bin/env python
"""Top SASS instructions of a kernel profile by stall samples, with the source line each maps to.
    python tools/ncu_top.py src.csv k.sass <kernel-substring> [N]"""
import csv, os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_by_line as n
src_csv, sass, func = sys.argv[1:4]
N = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
sl = n.sass_lines(sass, func)
base = int(data[0][0], 16)
by_off = {o: l for o, t, l in sl}
tot = sum(int(r[col["# Samples"]] or 0) for r in data)
last = None
ann = []
for r in data:
    l = by_off.get(int(r[0], 16) - base)
    if l: last = l
    ann.append((int(r[col["# Samples"]] or 0), int(r[col["Instructions Executed"]] or 0), r[col["Source"]].strip(), last, int(r[0], 16) - base))
for s, e, t, l, off in sorted(ann, key=lambda a: -a[0])[:N]:
    print(f"{100 * s / tot:5.2f}% {e:>11d} {off:6x} {t[:52]:52s} {l[0] + ':' + str(l[1]) if l else ''}")
