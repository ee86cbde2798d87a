"""Seeded synthetic genome sets for parity tests and bench.py.

Mutated families: a random root genome per family, members derived from the root by
substitutions (rate d), short indels (rate d/10, geometric length mean 5, cap 50), with
probability 0.3 one block event (inversion / translocation / duplication) and for 1% of the
members a run of 10-200 N.  Everything is driven by numpy's PCG64 with an explicit seed.
"""
from __future__ import annotations

import numpy as np

SEED = 20261104
_COMP = np.array([3, 2, 1, 0, 5, 5], dtype=np.uint8)


def _mutate(rng: np.random.Generator, root: np.ndarray, d: float, block=(500, 3000), p_block=0.3,
            p_nrun=0.01) -> np.ndarray:
    n = root.size
    seq = root.copy()
    # substitutions
    hit = rng.random(n) < d
    seq[hit] = (seq[hit] + rng.integers(1, 4, size=int(hit.sum()), dtype=np.uint8)) % 4
    # short indels
    n_indel = rng.binomial(n, d / 10.0)
    if n_indel:
        pos = np.sort(rng.integers(0, n, size=n_indel))
        lens = np.minimum(rng.geometric(1.0 / 5.0, size=n_indel), 50)
        is_ins = rng.random(n_indel) < 0.5
        parts = []
        prev = 0
        for p, l, ins in zip(pos, lens, is_ins):
            if p < prev:
                continue
            parts.append(seq[prev:p])
            if ins:
                parts.append(rng.integers(0, 4, size=int(l), dtype=np.uint8))
                prev = p
            else:
                prev = min(n, p + int(l))
        parts.append(seq[prev:])
        seq = np.concatenate(parts)
    # one block event
    if rng.random() < p_block and seq.size > 2 * block[1]:
        bl = int(rng.integers(block[0], block[1] + 1))
        a = int(rng.integers(0, seq.size - bl))
        kind = int(rng.integers(0, 3))
        blk = seq[a : a + bl].copy()
        if kind == 0:      # inversion (reverse complement in place)
            seq[a : a + bl] = _COMP[blk[::-1]]
        elif kind == 1:    # translocation
            rest = np.concatenate([seq[:a], seq[a + bl :]])
            b = int(rng.integers(0, rest.size))
            seq = np.concatenate([rest[:b], blk, rest[b:]])
        else:              # duplication
            b = int(rng.integers(0, seq.size))
            seq = np.concatenate([seq[:b], blk, seq[b:]])
    if rng.random() < p_nrun and seq.size > 400:
        l = int(rng.integers(10, 201))
        a = int(rng.integers(0, seq.size - l))
        seq[a : a + l] = 5
    return seq


def families(n_families: int, members: int, len_lo: int, len_hi: int, d_lo: float = 0.01, d_hi: float = 0.30,
             seed: int = SEED, block=(500, 3000)):
    """Returns (names, list of uint8 code arrays).  Member 0 of a family is mutated too."""
    rng = np.random.default_rng(seed)
    names, seqs = [], []
    for f in range(n_families):
        root = rng.integers(0, 4, size=int(rng.integers(len_lo, len_hi + 1)), dtype=np.uint8)
        for m in range(members):
            d = float(rng.uniform(d_lo, d_hi))
            seqs.append(_mutate(rng, root, d, block=block))
            names.append(f"f{f}_m{m}")
    return names, seqs


def viral(n_genomes: int, seed: int = SEED):
    """BASELINE.json configs[1]/[2] shape: families of 20, ~50 kb, 1-30 % divergence."""
    fam = max(1, (n_genomes + 19) // 20)
    names, seqs = families(fam, 20, 40_000, 60_000, seed=seed)
    return names[:n_genomes], seqs[:n_genomes]
