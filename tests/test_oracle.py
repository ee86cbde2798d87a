"""CPU tests (-m "not gpu"): pin oracle/lzani_oracle.c against the golden vectors generated from
the unmodified reference (tests/golden/make_golden.py) and, where oracle/_ref was built, against
the reference CParser live on fresh seeded inputs."""
import importlib
import os

import numpy as np
import pytest

from oracle_lib import Oracle, RefShim, hostio
import make_golden as mg

synth = importlib.import_module("lz-ani_b200.synth")
GOLD = np.load(os.path.join(os.path.dirname(mg.__file__), "golden.npz"))


@pytest.fixture(scope="module")
def oracle():
    return Oracle()


@pytest.mark.parametrize("case", list(mg.CASES))
def test_oracle_matches_reference_golden(oracle, case):
    sname, params = mg.CASES[case]
    ss = mg.get_set(sname)
    got = oracle.run_rows(ss, params)
    want = GOLD[case]
    assert got.shape == want.shape
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"{case}: {bad.size} pairs differ, first {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"


def test_vir61_checksums(oracle):
    # known answers recorded in BASELINE.md / SURVEY 8c for the reference's CI fixture
    assert GOLD["vir61_default"].sum(axis=0).tolist() == [36285147, 10563899, 332871]
    assert GOLD["vir61_alt"].sum(axis=0).tolist() == [25475944, 6034578, 88640]


@pytest.mark.parametrize("pair", [(0, 1), (1, 0), (5, 6), (7, 3), (10, 11)])
def test_oracle_regions_match_reference_golden(oracle, pair):
    ex = mg.get_set("example")
    reg, st = oracle.regions(ex, None, *pair)
    want = GOLD[f"example_regions_{pair[0]}_{pair[1]}"]
    assert np.array_equal(reg, want)
    # the kept regions' totals are the pair's calc_stats() whenever length and M+X agree
    n = ex and len(ex)
    k = pair[0] * (n - 1) + (pair[1] if pair[1] < pair[0] else pair[1] - 1)
    assert st.tolist() == GOLD["example_default"][k].tolist()


def test_sparse_rows_equal_dense_subset(oracle):
    ss = mg.get_set("synth_small")
    dense = oracle.run_rows(ss, None)
    n = len(ss)
    rows = np.array([7, 2, 9], dtype=np.uint32)
    off = np.array([0, 2, 2, 5], dtype=np.uint64)
    q = np.array([3, 0, 11, 1, 2], dtype=np.uint32)
    got = oracle.run_rows(ss, None, rows=rows, row_off=off, q_ids=q)
    k = 0
    for j, r in enumerate(rows):
        for e in range(int(off[j]), int(off[j + 1])):
            qq = int(q[e])
            want = dense[int(r) * (n - 1) + (qq if qq < r else qq - 1)]
            assert got[k].tolist() == want.tolist()
            k += 1


@pytest.mark.skipif(not RefShim.available(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("seed,params", [(1, None), (2, mg.ALT), (3, mg.ODD), (4, dict(mal=12, msl=8, reg=20, aw=32, am=10, ar=1))])
def test_oracle_vs_reference_live(oracle, seed, params):
    names, seqs = synth.families(2, 5, 2000, 6000, seed=seed, block=(100, 500))
    ss = hostio.SeqSet(names, seqs)
    a = RefShim().run_rows(ss, params)
    b = oracle.run_rows(ss, params)
    assert np.array_equal(a, b)
