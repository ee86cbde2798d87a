#!/usr/bin/env python
"""Throughput + parity spot checks of the BASELINE.json configs that are NOT the bench line
(configs[3] sparse phage list, configs[4] bacterial-scale genomes), scaled to one GPU and a few
minutes.  Prints one JSON line per config.  Parity: a sample of the pairs against the unmodified
reference CParser (oracle/_ref) when present, else the oracle port.

    python tools/other_configs.py [--phage-families 100] [--bact-genomes 8] [--bact-len 5000000]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
lz = importlib.import_module("lz-ani_b200")
hostio, synth = lz.hostio, lz.synth
from oracle_lib import Oracle, RefShim  # noqa: E402  (checker only)

LONGP = dict(mal=15, msl=9, mrd=80, mqd=80, reg=100, aw=20, am=9, ar=4)      # SURVEY 8d.5 long-genome params


def checker():
    return (RefShim(), "reference") if RefShim.available() else (Oracle(), "port")


def ordered(names, seqs):
    order = hostio.reorder(names, [s.size for s in seqs])
    return hostio.SeqSet([names[i] for i in order], [seqs[i] for i in order])


def phage_sparse(n_fam):
    """configs[3] shape: families of 20, 30-70 kb; pair list = all within-family pairs plus 5 %
    random cross-family decoys, as edges (i, j < i) of a kmer-db style filter; the CSR is built on
    the device (lzani_build_pair_list)."""
    names, seqs = synth.families(n_fam, 20, 30_000, 70_000, seed=synth.SEED + 4)
    ss = ordered(names, seqs)
    n = len(ss)
    fam = np.array([int(x.split("_")[0][1:]) for x in ss.names])
    rng = np.random.default_rng(4)
    ei, ej = np.nonzero(np.tril(fam[:, None] == fam[None, :], -1))
    n_decoy = max(1, ei.size // 20)
    di = rng.integers(1, n, n_decoy)
    dj = (rng.random(n_decoy) * di).astype(np.int64)
    keep = fam[di] != fam[dj]
    ei = np.concatenate([ei, di[keep]]).astype(np.uint32)
    ej = np.concatenate([ej, dj[keep]]).astype(np.uint32)
    key = np.unique(ei.astype(np.int64) * n + ej)
    ei, ej = (key // n).astype(np.uint32), (key % n).astype(np.uint32)
    m = lz.Matcher(device=0)
    m.set_sequences(ss)
    off, q = m.build_pair_list(n, ei, ej)
    pl_ms = m.timing()["pairlist_ms"]
    rows = np.arange(n, dtype=np.uint32)
    m.run_rows(rows, off, q, resident_repeat=1)
    t0 = time.perf_counter()
    res = m.run_rows(rows, off, q)
    wall = time.perf_counter() - t0
    t = m.timing()
    chk, kind = checker()
    srows = np.linspace(0, n - 1, 16).astype(np.uint32)
    soff, sq = lz.sharding.sub_csr(srows, off, q)
    want = chk.run_rows(ss, None, rows=srows, row_off=soff, q_ids=sq, threads=os.cpu_count())
    got = np.concatenate([res[int(off[r]):int(off[r + 1])] for r in srows])
    m.close()
    return {"config": "sparse phage list (BASELINE configs[3] shape)", "genomes": n, "ordered_pairs": int(off[-1]),
            "pairs_per_s_kernels": t["n_pairs"] / ((t["index_ms"] + t["parse_ms"]) / 1e3),
            "pairs_per_s_e2e": int(off[-1]) / wall, "query_gbp_per_s": t["q_bases"] / ((t["index_ms"] + t["parse_ms"]) / 1e3) / 1e9,
            "index_ms": t["index_ms"], "parse_ms": t["parse_ms"], "pairlist_ms": pl_ms,
            "cpu_pairs_per_s": len(want) / chk.last_seconds, "cpu_kind": kind, "cpu_cores": os.cpu_count(),
            "parity_sample_pairs": len(want), "parity": bool(np.array_equal(got, want))}


def bacterial(n_genomes, length):
    """configs[4] shape: families of 4, ~5 Mb, 1-15 % divergence, block events 5-50 kb, long-genome params."""
    fam = max(1, (n_genomes + 3) // 4)
    names, seqs = synth.families(fam, 4, int(length * 0.9), int(length * 1.1), d_lo=0.01, d_hi=0.15,
                                 seed=synth.SEED + 5, block=(5_000, 50_000))
    ss = ordered(names[:n_genomes], seqs[:n_genomes])
    n = len(ss)
    m = lz.Matcher(params=LONGP, device=0)
    m.set_sequences(ss)
    rows = np.arange(n, dtype=np.uint32)
    m.run_rows(rows, resident_repeat=1)      # warm-up at full size: the chunk scratch of the staged parse is allocated here
    t0 = time.perf_counter()
    res = m.run_rows(rows)
    wall = time.perf_counter() - t0
    t = m.timing()
    chk, kind = checker()
    want = chk.run_rows(ss, LONGP, rows=rows[:1], threads=os.cpu_count())
    m.close()
    return {"config": "bacterial-scale all2all (BASELINE configs[4] shape)", "genomes": n, "mean_len": float(ss.lens.mean()),
            "params": LONGP, "ordered_pairs": n * (n - 1),
            "pairs_per_s_kernels": t["n_pairs"] / ((t["index_ms"] + t["parse_ms"]) / 1e3), "pairs_per_s_e2e": n * (n - 1) / wall,
            "query_gbp_per_s": t["q_bases"] / ((t["index_ms"] + t["parse_ms"]) / 1e3) / 1e9,
            "index_ms": t["index_ms"], "parse_ms": t["parse_ms"],
            "cpu_pairs_per_s": len(want) / chk.last_seconds, "cpu_kind": kind, "cpu_cores": min(os.cpu_count(), n - 1),
            "parity_sample_pairs": len(want), "parity": bool(np.array_equal(res[: n - 1], want))}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--phage-families", type=int, default=100)
    ap.add_argument("--bact-genomes", type=int, default=8)
    ap.add_argument("--bact-len", type=int, default=5_000_000)
    a = ap.parse_args()
    if a.phage_families:
        print(json.dumps(phage_sparse(a.phage_families)), flush=True)
    if a.bact_genomes:
        print(json.dumps(bacterial(a.bact_genomes, a.bact_len)), flush=True)


if __name__ == "__main__":
    main()
