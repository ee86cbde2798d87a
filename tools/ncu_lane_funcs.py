#!/usr/bin/env python
"""Per-function breakdown of a k_walk_lanes profile (ncu --page source --csv + nvdisasm -g of the same cubin):
executed warp / thread instructions, stall samples and threads per instruction, by function of lane_walk.h and
by region of parse_core.h.

    python tools/ncu_lane_funcs.py src.csv k.sass [kernel-substring]
"""
import csv
import os
import re
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_by_line as n  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    src_csv, sass = sys.argv[1], sys.argv[2]
    func_sub = sys.argv[3] if len(sys.argv) > 3 else "k_walk_lanesINS_10PrmDefault"
    rows = list(csv.reader(open(src_csv)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
    sl = n.sass_lines(sass, func_sub)
    base = int(data[0][0], 16)
    by_off = {o: (t, l) for o, t, l in sl}
    marks = {}
    for fn in ("lane_walk.h", "parse_core.h", "kernels.cu", "layout.h"):
        src = open(os.path.join(ROOT, "lz-ani_b200", "csrc", fn)).read().split("\n")
        mk = []
        for i, l in enumerate(src):
            m = re.match(r"^(?:LZ_HD|LZ_COLD|inline|__device__ __forceinline__|__global__)?\s*[\w<>:&\*]+\s+(\w+)\(", l)
            if m and not l.startswith(" "):
                mk.append((i + 1, m.group(1)))
        marks[fn] = mk

    def func(fn, line):
        f = "head"
        for i, nm in marks.get(fn, []):
            if line >= i:
                f = nm
        return f

    agg = defaultdict(lambda: [0, 0, 0])
    tot = [0, 0, 0]
    stalls = defaultdict(lambda: defaultdict(int))
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for r in data:
        ent = by_off.get(int(r[0], 16) - base)
        if not ent or not ent[1]:
            key = "?"
        else:
            f, l = ent[1]
            key = f"{f.split('.')[0]}:{func(f, l)}" if f in marks else f
        ie = int(r[col["Instructions Executed"]] or 0)
        te = int(r[col["Thread Instructions Executed"]] or 0)
        sm = int(r[col["# Samples"]] or 0)
        a = agg[key]
        a[0] += ie; a[1] += te; a[2] += sm
        tot[0] += ie; tot[1] += te; tot[2] += sm
        for h in stall_cols:
            v = int(r[col[h]] or 0)
            if v:
                stalls[key][h[6:]] += v
    print(f"# warp-instr {tot[0]}  thread-instr {tot[1]}  thr/instr {tot[1] / max(tot[0], 1):.1f}  samples {tot[2]}")
    print(f"{'function':36s} {'winstr%':>8s} {'thread%':>8s} {'samples%':>8s} {'thr/instr':>9s}  top stalls")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][2]):
        if a[0] * 500 < tot[0] and a[2] * 500 < tot[2]:
            continue
        st = " ".join(f"{x}={v}" for x, v in sorted(stalls[k].items(), key=lambda kv: -kv[1])[:3])
        print(f"{k:36s} {100 * a[0] / tot[0]:8.2f} {100 * a[1] / tot[1]:8.2f} {100 * a[2] / tot[2]:8.2f} {a[1] / max(a[0], 1):9.1f}  {st}")


if __name__ == "__main__":
    main()
