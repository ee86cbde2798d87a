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


@pytest.mark.skipif(not RefShim.available(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("seed", list(range(200000, 200048)))
def test_oracle_vs_reference_random_parameters(oracle, seed):
    """The oracle against the UNMODIFIED reference CParser on random parameter sets and ragged low-complexity sets (the
    generators of tests/test_emul.py).  mqd is kept <= mrd: beyond that the reference reads seq_ref past its end
    (compare_ranges_both_ways, parser.cpp:288, has no bound; a literal run of up to mqd positions behind the last
    match may reach further than the mrd padding symbols at the end of seq_ref), so its own results depend on what the
    heap holds there.  Offline: 4 325 sets with mqd <= mrd, no difference; 1 675 sets with mqd > mrd, 19 differ."""
    from test_emul import _weird_sequence
    rng = np.random.default_rng(seed)
    if seed % 3 == 0:
        params = None
    elif seed % 3 == 1:
        params = dict(mal=11, msl=7, mrd=int(rng.integers(1, 110)), mqd=int(rng.integers(0, 70)), reg=int(rng.integers(1, 80)),
                      aw=15, am=7, ar=3)
    else:
        mal = int(rng.integers(3, 19))
        params = dict(mal=mal, msl=int(rng.integers(2, min(mal + 3, 13))), mrd=int(rng.integers(1, 120)), mqd=int(rng.integers(0, 130)),
                      reg=int(rng.integers(1, 90)), aw=int(rng.integers(1, 33)), am=int(rng.integers(0, 14)), ar=int(rng.integers(1, 7)))
    if params is not None and params["mqd"] > params["mrd"]:
        params["mqd"] = params["mrd"]
    if seed % 2:
        base = rng.integers(0, 4, int(rng.integers(50, 3000)), dtype=np.uint8)
        seqs = [base] + [_weird_sequence(rng, base) for _ in range(int(rng.integers(2, 5)))]
        if seed % 7 == 0:
            seqs.append(np.zeros(int(rng.integers(0, 12)), np.uint8))
        names = [f"s{i}" for i in range(len(seqs))]
    else:
        names, seqs = synth.families(2, 4, 1500, 5000, seed=seed, d_lo=0.0 if seed % 4 == 0 else 0.02,
                                     d_hi=[0.06, 0.2, 0.4][seed % 3], block=(60, 300))
        seqs = [s.copy() for s in seqs]
        a = int(rng.integers(0, 900))
        seqs[2][a:a + int(rng.integers(1, 60))] = 4
    ss = hostio.SeqSet(list(names), list(seqs))
    assert np.array_equal(oracle.run_rows(ss, params, threads=4), RefShim().run_rows(ss, params, threads=4)), params
