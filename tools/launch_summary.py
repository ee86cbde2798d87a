#!/usr/bin/env python
"""Aggregate an ncu launch list (--csv, one metric per row) by kernel: launches, total ms, warp instructions,
issue-slot utilisation, occupancy, registers, threads per instruction."""
import collections
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ik, im, iv = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value')
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[1:]:
    agg[r[ik].split('(')[0].replace('void ', '').replace('lzani::', '')[:44]][r[im]].append(float(r[iv].replace(',', '')))
tot = sum(sum(m['gpu__time_duration.sum']) for m in agg.values())
print(f"{'kernel':46s} {'n':>5s} {'total_ms':>9s} {'share':>6s} {'avg_us':>9s} {'Ginst':>8s} {'issue%':>6s} {'warps%':>6s} {'regs':>4s} {'thr/i':>5s}")
for name, m in agg.items():
    t = m['gpu__time_duration.sum']
    n = len(t)
    g = lambda k: sum(m[k]) / n if m.get(k) else float('nan')
    print(f"{name:46s} {n:5d} {sum(t)/1e6:9.2f} {100*sum(t)/tot:5.1f}% {sum(t)/n/1e3:9.1f} {sum(m.get('smsp__inst_executed.sum',[0]))/1e9:8.2f} "
          f"{g('smsp__issue_active.avg.pct_of_peak_sustained_active'):6.1f} {g('sm__warps_active.avg.pct_of_peak_sustained_active'):6.1f} "
          f"{g('launch__registers_per_thread'):4.0f} {g('smsp__thread_inst_executed_per_inst_executed.ratio'):5.1f}")
print(f"{'sum (serialised under ncu)':46s} {'':5s} {tot/1e6:9.2f}")
