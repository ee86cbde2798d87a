#!/usr/bin/env python
"""Generates the committed golden vectors from the UNMODIFIED reference (oracle/_ref, built by
`make -C oracle ref` from /root/reference).  Run in the build container only:

    python tests/golden/make_golden.py

Outputs (all small, committed):
  vir61.npz      the reference's CI fixture test/vir61 as three_in_byte bytes (post-reorder ids)
  example.npz    example/multifasta.fna likewise
  golden.npz     per case: int32 (n_pairs, 3) results of the reference CParser in all2all CSR order
                 (+ regions of a few pairs for the --out-alignment path)
  vir61.ani.tsv.md5 etc.: checksums of the reference's golden TSVs, for the record
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import RefShim, hostio, ROOT  # noqa: E402
import importlib  # noqa: E402

synth = importlib.import_module("lz-ani_b200.synth")
REF = "/root/reference"

ALT = dict(mal=15, msl=9, mrd=60, mqd=60, reg=60, aw=20, am=9, ar=4)
ODD = dict(mal=9, msl=5, mrd=10, mqd=60, reg=20, aw=8, am=3, ar=2)       # mqd > mrd: wrapped exponent
LONGP = dict(mal=15, msl=9, mrd=80, mqd=80, reg=100, aw=20, am=9, ar=4)  # long-genome params

CASES = {
    # name: (set, params)
    "vir61_default": ("vir61", None),
    "vir61_alt": ("vir61", ALT),
    "example_default": ("example", None),
    "example_alt": ("example", ALT),
    "example_odd": ("example", ODD),
    "synth_small_default": ("synth_small", None),
    "synth_small_long": ("synth_small", LONGP),
    "edge_default": ("edge", None),
    "edge_odd": ("edge", ODD),
}


def save_set(path, ss):
    np.savez_compressed(path, names=np.array(ss.names), lens=ss.lens,
                        packed=np.concatenate(ss.packed) if len(ss) else np.zeros(0, np.uint8))


def load_set(path):
    z = np.load(path)
    lens = z["lens"]
    packed = z["packed"]
    codes, o = [], 0
    for n in lens:
        nb = (int(n) + 2) // 3
        codes.append(hostio.unpack3(packed[o:o + nb], int(n)))
        o += nb
    return hostio.SeqSet([str(x) for x in z["names"]], codes)


def synth_small():
    names, seqs = synth.families(3, 4, 3000, 5000, seed=synth.SEED, block=(200, 600))
    return hostio.SeqSet.from_records([(n, bytes(np.array(list(b"ACGTNN"), np.uint8)[s])) for n, s in zip(names, seqs)])


def edge_set():
    """Tiny / degenerate inputs: empty, shorter than a k-mer, all-N, homopolymer, N-rich,
    exact duplicate, reverse complement, tandem repeat."""
    rng = np.random.default_rng(7)
    base = rng.integers(0, 4, size=900, dtype=np.uint8)
    comp = np.array([3, 2, 1, 0], np.uint8)
    nrich = base.copy()
    nrich[rng.random(base.size) < 0.05] = 5
    unit = rng.integers(0, 4, size=13, dtype=np.uint8)
    items = [
        ("empty", np.zeros(0, np.uint8)),
        ("tiny5", base[:5]),
        ("tiny12", base[:12]),
        ("allN", np.full(50, 5, np.uint8)),
        ("polyA", np.zeros(300, np.uint8)),
        ("base", base),
        ("dup", base.copy()),
        ("revcomp", comp[base[::-1]]),
        ("nrich", nrich),
        ("tandem", np.tile(unit, 40)),
        ("prefix", base[:400]),
        ("shifted", np.concatenate([rng.integers(0, 4, size=37, dtype=np.uint8), base[100:]])),
    ]
    return hostio.SeqSet([n for n, _ in items], [c for _, c in items])


def get_set(name):
    if name == "vir61":
        return load_set(os.path.join(HERE, "vir61.npz"))
    if name == "example":
        return load_set(os.path.join(HERE, "example.npz"))
    if name == "synth_small":
        return synth_small()
    if name == "edge":
        return edge_set()
    raise KeyError(name)


def main():
    ref = RefShim()
    save_set(os.path.join(HERE, "vir61.npz"), hostio.SeqSet.from_records(hostio.read_fasta_dir(f"{REF}/test/vir61")))
    save_set(os.path.join(HERE, "example.npz"), hostio.SeqSet.from_records(hostio.read_fasta(f"{REF}/example/multifasta.fna")))
    out = {}
    for case, (sname, params) in CASES.items():
        ss = get_set(sname)
        res = ref.run_rows(ss, params)
        out[case] = res
        print(case, len(ss), res.shape, res.sum(0))
    # regions of a few pairs (reverse-strand and rearranged ones included)
    ex = get_set("example")
    for (r, q) in [(0, 1), (1, 0), (5, 6), (7, 3), (10, 11)]:
        reg, st = ref.regions(ex, None, r, q)
        out[f"example_regions_{r}_{q}"] = reg
    np.savez_compressed(os.path.join(HERE, "golden.npz"), **out)
    with open(os.path.join(HERE, "reference_md5.txt"), "w") as f:
        for p in ["test/vir61.ani.tsv", "test/vir61.ani.ids.tsv", "example/output/ani.tsv"]:
            f.write(hashlib.md5(open(f"{REF}/{p}", "rb").read()).hexdigest() + "  " + p + "\n")


if __name__ == "__main__":
    main()
