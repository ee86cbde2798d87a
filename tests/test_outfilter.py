"""--out-filter on the device (reference src/lz_matcher.cpp:404-463).

The golden TSVs tests/golden/*outfilter*.ani.tsv.gz were written by the UNMODIFIED reference CLI with
`--out-filter` (tests/golden/make_golden_tsv.py).  A line (qidx, ridx) is in such a file iff the
reference's double arithmetic let it pass, so the set of lines pins the decision function:

  * CPU (`-m "not gpu"`): the decision function of the F1 kernel (csrc/outfilter.h), compiled for
    the host by tests/emul and fed the reference CParser's integers (golden.npz), must select
    exactly the golden lines; and it must agree with a numpy float64 restatement on random integers.
  * GPU (`-m gpu`): lzani_filter_pairs through the C ABI must return exactly those pairs / line bits,
    for the dense list, a sparse symmetric CSR list and a row-sharded group of two contexts.
"""
import ctypes as C
import gzip
import importlib
import os

import numpy as np
import pytest

from oracle_lib import Emul
import make_golden as mg

lz = importlib.import_module("lz-ani_b200")
GOLD_DIR = os.path.dirname(mg.__file__)
GOLD = np.load(os.path.join(GOLD_DIR, "golden.npz"))
KEYS = ("gani", "ani", "tani", "qcov", "rcov")

# case -> (set, golden results key, thresholds)
DENSE_CASES = {
    "vir61_outfilter": ("vir61", "vir61_default", dict(ani=0.9, qcov=0.3)),
    "example_outfilter": ("example", "example_default", dict(tani=0.85, rcov=0.9, gani=0.8)),
}


def golden_lines(case):
    """{(ridx, qidx)} of the reference CLI's filtered TSV (columns qidx ridx ...)."""
    txt = gzip.open(os.path.join(GOLD_DIR, case + ".ani.tsv.gz")).read().decode().splitlines()
    assert txt[0].split("\t")[:2] == ["qidx", "ridx"]
    return {(int(l.split("\t")[1]), int(l.split("\t")[0])) for l in txt[1:]}


def dense_pairs(n, res):
    """all2all CSR results (n*(n-1), 3) -> arrays over unordered pairs a < b: a, b, ab, ba."""
    a, b = np.triu_indices(n, 1)
    ab = res[a * (n - 1) + (b - 1)]          # row a, query b > a
    ba = res[b * (n - 1) + a]                # row b, query a < b
    return a.astype(np.uint32), b.astype(np.uint32), np.ascontiguousarray(ab), np.ascontiguousarray(ba)


def emul_masks(emul, thr, ab, ba, len_a, len_b):
    t = np.array([float(thr.get(k, 0.0)) for k in KEYS], dtype=np.float64)
    m = np.zeros(len(ab), dtype=np.uint32)
    la = np.ascontiguousarray(len_a, dtype=np.uint32)
    lb = np.ascontiguousarray(len_b, dtype=np.uint32)
    emul.lib.lzani_emul_out_filter(t.ctypes.data_as(C.c_void_p), C.c_uint64(len(ab)), ab.ctypes.data_as(C.c_void_p),
                                   ba.ctypes.data_as(C.c_void_p), la.ctypes.data_as(C.c_void_p),
                                   lb.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p))
    return m


def lines_of(a, b, masks):
    s = set()
    for x, y, m in zip(a.tolist(), b.tolist(), masks.tolist()):
        if m & 1:
            s.add((x, y))       # line 0: reference a, query b
        if m & 2:
            s.add((y, x))       # line 1: reference b, query a
    return s


def numpy_masks(thr, ab, ba, len_a, len_b):
    """float64 restatement of lz_matcher.cpp:440-462 (numpy division is IEEE double)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        ln = [len_b.astype(np.uint32), len_a.astype(np.uint32)]
        mat = [ab[:, 0], ba[:, 0]]
        lit = [ab[:, 1], ba[:, 1]]
        tani = (mat[0] + mat[1]).astype(np.float64) / (ln[0] + ln[1]).astype(np.float64)
        gani = [mat[i].astype(np.float64) / ln[i].astype(np.float64) for i in range(2)]
        den = [(mat[i] + lit[i]) for i in range(2)]
        ani = [np.where(den[i] != 0, mat[i].astype(np.float64) / np.where(den[i] != 0, den[i], 1).astype(np.float64), 0.0) for i in range(2)]
        cov = [den[i].astype(np.float64) / ln[i].astype(np.float64) for i in range(2)]
        g = lambda k: float(thr.get(k, 0.0))
        m = np.zeros(len(ab), np.uint32)
        for i in range(2):
            ok = ~(gani[i] < g("gani")) & ~(ani[i] < g("ani")) & ~(tani < g("tani")) & ~(cov[i] < g("qcov")) & ~(cov[1 - i] < g("rcov"))
            m |= ok.astype(np.uint32) << i
    return m


@pytest.fixture(scope="module")
def emul():
    return Emul()


@pytest.mark.parametrize("case", list(DENSE_CASES))
def test_decision_function_selects_the_reference_lines(emul, case):
    sname, gkey, thr = DENSE_CASES[case]
    ss = mg.get_set(sname)
    n = len(ss)
    a, b, ab, ba = dense_pairs(n, GOLD[gkey])
    masks = emul_masks(emul, thr, ab, ba, ss.lens[a], ss.lens[b])
    want = golden_lines(case)
    assert 0 < len(want) < n * (n - 1)
    assert lines_of(a, b, masks) == want
    assert np.array_equal(masks, numpy_masks(thr, ab, ba, ss.lens[a], ss.lens[b]))


def test_decision_function_on_random_integers_and_edges(emul):
    rng = np.random.default_rng(99)
    n = 200000
    la = rng.integers(0, 60000, n).astype(np.uint32)
    lb = rng.integers(0, 60000, n).astype(np.uint32)
    ab = np.stack([rng.integers(0, 60000, n), rng.integers(0, 20000, n), rng.integers(0, 300, n)], 1).astype(np.int32)
    ba = np.stack([rng.integers(0, 60000, n), rng.integers(0, 20000, n), rng.integers(0, 300, n)], 1).astype(np.int32)
    ab[:1000, :2] = 0                      # mat + lit == 0: ani defined as 0
    la[500:1500] = 0                       # zero lengths: 0/0 = NaN and x/0 = inf never compare below
    for thr in (dict(ani=0.7), dict(gani=0.5, qcov=0.8), dict(tani=0.4, rcov=0.6), dict(ani=1.0, gani=1.0), {},
                dict(ani=0.33333333333333331, qcov=2.0 / 3.0)):
        assert np.array_equal(emul_masks(emul, thr, ab, ba, la, lb), numpy_masks(thr, ab, ba, la, lb)), thr
    # exact ties: a metric equal to its threshold passes (the reference tests `<`)
    ab2 = np.array([[3, 1, 1]], np.int32); ba2 = np.array([[3, 1, 1]], np.int32)
    l4 = np.array([4], np.uint32)
    assert emul_masks(emul, dict(ani=0.75, gani=0.75, tani=0.75, qcov=1.0, rcov=1.0), ab2, ba2, l4, l4)[0] == 3
    assert emul_masks(emul, dict(ani=0.7500000000000001), ab2, ba2, l4, l4)[0] == 0


# ---------------------------------------------------------------------------------------- GPU
def run_keep(m, rows, row_off=None, q_ids=None):
    m.run_rows(rows, row_off, q_ids, resident_repeat=1)


@pytest.mark.gpu
@pytest.mark.parametrize("case", list(DENSE_CASES))
def test_filter_pairs_dense_matches_reference_cli(case):
    sname, gkey, thr = DENSE_CASES[case]
    ss = mg.get_set(sname)
    n = len(ss)
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        run_keep(m, np.arange(n, dtype=np.uint32))
        got = m.filter_pairs(ss.lens, thr, cap=4)       # tiny cap: exercises the NOMEM retry
        t = m.timing()
    finally:
        m.close()
    assert lines_of(got["a"], got["b"], got["lines"]) == golden_lines(case)
    assert (got["lines"] != 0).all() and (got["a"] < got["b"]).all()
    # CSR order, and the integers are the reference's
    key = got["a"].astype(np.int64) * n + got["b"]
    assert (np.diff(key) > 0).all()
    res = GOLD[gkey]
    assert np.array_equal(got["ab"], res[got["a"].astype(np.int64) * (n - 1) + got["b"] - 1])
    assert np.array_equal(got["ba"], res[got["b"].astype(np.int64) * (n - 1) + got["a"]])
    assert t["filter_ms"] > 0 and t["filter_d2h_bytes"] < 12 * n * (n - 1)


def symmetric_sparse_list(n, rng, density=0.3):
    keep = np.triu(rng.random((n, n)) < density, 1)
    keep = keep | keep.T
    off, q = [0], []
    for r in range(n):
        ids = np.nonzero(keep[r])[0]
        q.extend(int(x) for x in ids)
        off.append(len(q))
    return np.array(off, np.uint64), np.array(q, np.uint32), keep


@pytest.mark.gpu
def test_filter_pairs_sparse_and_two_context_group():
    """A symmetric sparse CSR list, run (1) in one context and (2) dealt r mod 2 over two contexts
    that read each other's rows in place (same device here; over NVLink peer access on two)."""
    ss = mg.get_set("example")
    n = len(ss)
    res = GOLD["example_default"]
    thr = dict(ani=0.9, qcov=0.5)
    off, q, keep = symmetric_sparse_list(n, np.random.default_rng(3))
    a, b = np.nonzero(np.triu(keep, 1))
    ab = res[a * (n - 1) + b - 1]
    ba = res[b * (n - 1) + a]
    want_masks = numpy_masks(thr, ab, ba, ss.lens[a], ss.lens[b])
    want = lines_of(a, b, want_masks)
    assert want

    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        run_keep(m, np.arange(n, dtype=np.uint32), off, q)
        got = m.filter_pairs(ss.lens, thr)
        assert lines_of(got["a"], got["b"], got["lines"]) == want
        # an asymmetric list is an error, as the reference's assert (lz_matcher.cpp:417)
        q2 = q.copy()
        off2 = off.copy()
        r0 = int(np.nonzero(np.diff(off.astype(np.int64)) > 0)[0][-1])     # drop the last entry of the last non-empty row
        q2 = np.delete(q2, int(off[r0 + 1]) - 1)
        off2[r0 + 1:] -= 1
        run_keep(m, np.arange(n, dtype=np.uint32), off2, q2)
        with pytest.raises(lz.LzaniError, match="not symmetric"):
            m.filter_pairs(ss.lens, thr)
    finally:
        m.close()

    import torch
    second = 1 if torch.cuda.device_count() >= 2 else 0      # two GPUs when the box has them: real peer access
    ms = [lz.Matcher(device=0), lz.Matcher(device=second)]
    try:
        parts = []
        for g, mm in enumerate(ms):
            mm.set_sequences(ss)
            rows = np.arange(g, n, 2, dtype=np.uint32)
            sub_off, sub_q = lz.sharding.sub_csr(rows, off, q)
            run_keep(mm, rows, sub_off, sub_q)
        for mm in ms:
            parts.append(mm.filter_pairs(ss.lens, thr, group=ms))
        got = np.concatenate(parts)
        assert lines_of(got["a"], got["b"], got["lines"]) == want
        assert len(got) == int((want_masks != 0).sum())
        # dense rows dealt over the two contexts
        for g, mm in enumerate(ms):
            run_keep(mm, np.arange(g, n, 2, dtype=np.uint32))
        got = np.concatenate([mm.filter_pairs(ss.lens, DENSE_CASES["example_outfilter"][2], group=ms) for mm in ms])
        assert lines_of(got["a"], got["b"], got["lines"]) == golden_lines("example_outfilter")
    finally:
        for mm in ms:
            mm.close()
