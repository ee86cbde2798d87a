#!/usr/bin/env python
"""Join an `ncu --page source --csv` dump (SASS level) with `nvdisasm -g` line info of the same
cubin and aggregate executed warp instructions / stall samples per CUDA source line.

    cuobjdump -xelf all lz-ani_b200/liblzani_b200.so      # -> kernels.sm_100a.cubin
    nvdisasm -g -c kernels.sm_100a.cubin > k.sass
    ncu -i prof.ncu-rep --page source --csv > src.csv
    python tools/ncu_by_line.py src.csv k.sass <mangled-kernel-substring> [top]
"""
import csv
import re
import sys
from collections import defaultdict


def sass_lines(path, func_sub):
    out, cur, on = [], None, False
    for ln in open(path):
        if ln.startswith(".text."):
            on = func_sub in ln
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out.append((int(m.group(1), 16), m.group(2).strip(), cur))
    return out


def main():
    src_csv, sass, func = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    sl = sass_lines(sass, func)
    base = int(data[0][0], 16)
    by_off = {o: (t, l) for o, t, l in sl}
    agg = defaultdict(lambda: [0, 0, 0, 0])
    tot_i = tot_s = 0
    miss = 0
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    stall_by_line = defaultdict(lambda: defaultdict(int))
    for r in data:
        off = int(r[0], 16) - base
        ent = by_off.get(off)
        if ent is None:
            miss += 1
            continue
        line = ent[1]
        ie = int(r[col["Instructions Executed"]] or 0)
        te = int(r[col["Thread Instructions Executed"]] or 0)
        sm = int(r[col["# Samples"]] or 0)
        a = agg[line]
        a[0] += ie; a[1] += te; a[2] += sm; a[3] += 1
        tot_i += ie; tot_s += sm
        for h in stall_cols:
            v = int(r[col[h]] or 0)
            if v:
                stall_by_line[line][h] += v
    print(f"# instructions matched {len(data) - miss}/{len(data)}; warp-instr executed {tot_i}; samples {tot_s}")
    print(f"{'file:line':28s} {'winstr%':>8s} {'samples%':>8s} {'thr/instr':>9s} {'sass':>5s}  top stalls")
    for line, a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        st = sorted(stall_by_line[line].items(), key=lambda kv: -kv[1])[:3]
        ss = " ".join(f"{k[6:]}={v}" for k, v in st)
        nm = f"{line[0]}:{line[1]}" if line else "?"
        print(f"{nm:28s} {100.0 * a[0] / max(tot_i, 1):8.2f} {100.0 * a[2] / max(tot_s, 1):8.2f} {a[1] / max(a[0], 1):9.1f} {a[3]:5d}  {ss}")


if __name__ == "__main__":
    main()
