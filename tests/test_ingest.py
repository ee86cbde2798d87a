"""Device-side ingest (reference CSeqReservoir::append, src/seq_reservoir.cpp:18-87):
lzani_set_sequences_text takes the loader's text (one byte per base) and the K1 kernel builds the
bit-plane records and the reverse complement from it, instead of from base-6 triples packed on the host.

  * CPU: K1's element functions (layout.h, compiled for the host by tests/emul) must give identical
    records from the text and from the reference's three_in_byte packing of the same text -- for
    upper/lower case, N, IUPAC codes, the byte-5 contig separators of load_fasta and ragged lengths.
  * GPU: parsing sets handed over as text gives the reference's golden integers.
"""
import ctypes as C
import importlib
import os

import numpy as np
import pytest

from oracle_lib import Emul, Oracle, hostio
import make_golden as mg

lz = importlib.import_module("lz-ani_b200")
GOLD = np.load(os.path.join(os.path.dirname(mg.__file__), "golden.npz"))
LETTERS = np.frombuffer(b"ACGTNN", dtype=np.uint8)


def messy_text(rng, n):
    """Loader text with everything dna_code (seq_reservoir.h:243-247) has to map."""
    t = LETTERS[rng.integers(0, 4, n)].copy()
    low = rng.random(n) < 0.3
    t[low] |= 0x20                                        # lower case
    amb = rng.random(n) < 0.02
    t[amb] = np.frombuffer(b"NnRYKMSWBDHVryu-*", dtype=np.uint8)[rng.integers(0, 17, int(amb.sum()))]
    if n > 200:
        t[100:140] = 5                                    # load_fasta's contig separator: bytes of value code_N_seq
    return t


def records(emul, src, length, mrd, ascii_):
    nq, nr = emul.lib.lzani_emul_q_recs(length, mrd), emul.lib.lzani_emul_r_recs(length, mrd)
    q = np.zeros((nq, 4), np.uint32)
    r = np.zeros((nr, 4), np.uint32)
    src = np.ascontiguousarray(src, dtype=np.uint8)
    buf = np.concatenate([src, np.zeros(1, np.uint8)])    # non-empty buffer for len 0
    emul.lib.lzani_emul_pack_records(buf.ctypes.data_as(C.c_void_p), C.c_uint32(length), C.c_int(mrd), C.c_int(ascii_),
                                     q.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p))
    return q, r


@pytest.mark.parametrize("n", [0, 1, 2, 3, 31, 32, 33, 64, 95, 1000, 4099])
@pytest.mark.parametrize("mrd", [7, 40])
def test_records_from_text_equal_records_from_reservoir_bytes(n, mrd):
    emul = Emul()
    t = messy_text(np.random.default_rng(n * 131 + mrd), n)
    packed = hostio.pack3(hostio.encode(t))               # what CSeqReservoir::append stores
    q3, r3 = records(emul, packed, n, mrd, 0)
    qa, ra = records(emul, t, n, mrd, 1)
    assert np.array_equal(q3, qa) and np.array_equal(r3, ra)
    # and the planes say what they should: N mask exactly where the text is not ACGTacgt
    codes = hostio.encode(t)
    nmask = np.unpackbits(qa[:, 2].astype("<u4").view(np.uint8), bitorder="little")[:n]
    assert np.array_equal(nmask.astype(bool), codes > 3)


def text_of(ss, rng=None):
    out = []
    for c in ss.codes:
        t = LETTERS[c].copy()
        if rng is not None and t.size:
            t[rng.random(t.size) < 0.5] |= 0x20
        out.append(t)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["example_default", "edge_default", "edge_odd", "vir61_alt"])
def test_text_ingest_gives_reference_golden(case):
    sname, params = mg.CASES[case]
    ss = mg.get_set(sname)
    m = lz.Matcher(params=params, device=0)
    try:
        m.set_sequences_text(text_of(ss, np.random.default_rng(1)))
        got = m.all2all()
        t = m.timing()
    finally:
        m.close()
    assert np.array_equal(got, GOLD[case])
    assert t["n_pairs"] == len(got)


@pytest.mark.gpu
def test_text_ingest_with_separators_and_iupac_vs_oracle():
    rng = np.random.default_rng(17)
    names, seqs = lz.synth.families(2, 4, 3000, 6000, seed=41, block=(200, 700))
    texts = []
    for s in seqs:
        t = LETTERS[s].copy()
        t[rng.random(t.size) < 0.4] |= 0x20
        pos = int(rng.integers(500, t.size - 500))
        t = np.concatenate([t[:pos], np.full(40, 5, np.uint8), t[pos:]])       # a two-contig genome (load_fasta :110-127)
        amb = rng.random(t.size) < 0.003
        t[amb] = np.frombuffer(b"RYKMSWn", dtype=np.uint8)[rng.integers(0, 7, int(amb.sum()))]
        texts.append(t)
    ss = hostio.SeqSet(names, [hostio.encode(t) for t in texts])
    m = lz.Matcher(device=0)
    try:
        m.set_sequences_text(texts)
        got = m.all2all()
    finally:
        m.close()
    assert np.array_equal(got, Oracle().run_rows(ss, None))
