#!/usr/bin/env python
"""Per-function-region breakdown of a kernel profile: executed warp instructions, stall samples and
code size, attributing inlined intrinsic lines to the enclosing parse_core.h / kernels.cu region.

    python tools/ncu_regions.py src.csv k.sass <kernel-substring> <pairs-per-launch>
"""
import csv
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_by_line as n  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MARKS = [("equal_len_coop", "int equal_len_coop("), ("approx_extend", "void approx_extend("),
         ("both_ways", "int both_ways_best("), ("anchor_lookup", "void anchor_lookup("),
         ("diag_helpers", "struct DiagWord"), ("extend_in_word", "struct ExtWord"), ("spec_first_seed", "struct V3"),
         ("hit_probe", "void hit_probe("), ("hit_extend", "void hit_extend("), ("take_match", "void take_match("), ("diag_seed_starts", "uint64_t diag_seed_starts("),
         ("best_seed", "void best_seed("), ("pp:init", "PairResult parse_pair("), ("pp:window_scan", "// ---- new window"),
         ("pp:batch_list", "// ---- new batch"), ("pp:batch_eval", "b_cover = total <= 32"), ("pp:batch_spec", "if (PT::kSpecSeeds && !RS::kActive && P.mqd <= SPEC_MAX_K)"), ("pp:fast", "// ---- FAST PATH"), ("pp:stored_load", "// ---- next batch: the first <= 32 stored"), ("pp:inline_batch", "} else {\n            if (i0 >= win_end)"), ("pp:chain", "// ---- chain: verified successor links"), ("pp:sK1", "// what the speculative seed record says"),
         ("scan_word", "LZ_HD uint32_t scan_word("), ("scan_coarse", "LZ_HD uint32_t scan_word_coarse("), ("hit_spec", "void hit_spec("), ("hit_record", "LZ_HD Rec hit_record("), ("verified_link", "LZ_HD uint32_t verified_link("), ("next_event", "struct Event {"), ("pp:nopred", "// ---- no prediction"),
         ("pp:pred", "// ---- prediction mode"), ("pp:take", "if (matched) {"), ("pp:tail", "// tail (:710-713)")]


def main():
    src_csv, sass, func, pairs = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
    src = open(os.path.join(ROOT, "lz-ani_b200", "csrc", "parse_core.h")).read().split("\n")
    starts = []
    for nm, pat in MARKS:
        hit = [i for i, x in enumerate(src) if pat in x]
        if hit:
            starts.append((hit[0] + 1, nm))
    starts.sort()

    def region(line):
        r = "head"
        for i, nm in starts:
            if line >= i:
                r = nm
        return r

    rows = list(csv.reader(open(src_csv)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    sl = n.sass_lines(sass, func)
    base = int(data[0][0], 16)
    by_off = {o: l for o, t, l in sl}
    agg = defaultdict(lambda: [0, 0, 0, 0, 0])
    last = "?"
    for r in data:
        l = by_off.get(int(r[0], 16) - base)
        if l and l[0] == "parse_core.h":
            last = region(l[1])
        elif l and l[0] == "kernels.cu":
            last = "kernels.cu"
        a = agg[last]
        a[0] += int(r[col["Instructions Executed"]] or 0)
        a[1] += int(r[col["# Samples"]] or 0)
        a[2] += 1
        a[3] += int(r[col["Thread Instructions Executed"]] or 0)
        a[4] += int(r[col.get("stall_no_inst", 0)] or 0) if "stall_no_inst" in col else 0
    ti = sum(a[0] for a in agg.values())
    ts = sum(a[1] for a in agg.values())
    print(f"# {ti} warp-instr, {ti / pairs:.0f} per pair; {len(data)} SASS instr")
    print(f"{'region':18s} {'winstr%':>8s} {'samples%':>9s} {'winstr/pair':>12s} {'thr/instr':>9s} {'sass':>6s} {'no_inst%':>8s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:18s} {100 * a[0] / ti:8.2f} {100 * a[1] / ts:9.2f} {a[0] / pairs:12.0f} {a[3] / max(a[0], 1):9.1f} {a[2]:6d} {100 * a[4] / max(a[1], 1):8.1f}")


if __name__ == "__main__":
    main()
