#!/usr/bin/env python
"""Profiling helper: one all2all of a seeded synthetic set through the library (no checks, no CPU legs).

    python tools/run_set.py <n_families> <members> [repeat]
members = 1 gives a set of unrelated genomes (every pair is the plain unrelated-pair case)."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lz = importlib.import_module("lz-ani_b200")

nf, mem = int(sys.argv[1]), int(sys.argv[2])
rep = int(sys.argv[3]) if len(sys.argv) > 3 else 1
names, seqs = lz.synth.families(nf, mem, 40_000, 60_000, seed=lz.synth.SEED)
ss = lz.hostio.SeqSet(names, seqs)
m = lz.Matcher(device=0)
m.set_sequences(ss)
for _ in range(rep):
    res = m.all2all()
    t = m.timing()
    print(f"{len(ss)} genomes, {res.shape[0]} pairs: parse {t['parse_ms']:.2f} ms = {res.shape[0] / t['parse_ms'] * 1e3:.0f} pairs/s, "
          f"index {t['index_ms']:.2f} ms, checksum {res.sum(axis=0).tolist()}")
