"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the golden
vectors generated from the unmodified reference and against the oracle on fresh inputs."""
import importlib
import os

import numpy as np
import pytest

from oracle_lib import Oracle, hostio
import make_golden as mg

lz = importlib.import_module("lz-ani_b200")
synth = lz.synth
GOLD = np.load(os.path.join(os.path.dirname(mg.__file__), "golden.npz"))
pytestmark = pytest.mark.gpu


def run_gpu(ss, params, **kw):
    m = lz.Matcher(params=params, device=0)
    try:
        m.set_sequences(ss)
        return m.all2all(), m.timing()
    finally:
        m.close()


@pytest.mark.parametrize("case", list(mg.CASES))
def test_cuda_matches_reference_golden(case):
    sname, params = mg.CASES[case]
    ss = mg.get_set(sname)
    got, _ = run_gpu(ss, params)
    want = GOLD[case]
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"{case}: {bad.size}/{len(want)} pairs differ, first {bad[:5]}: {got[bad[:3]]} vs {want[bad[:3]]}"


@pytest.mark.parametrize("seed,params", [
    (31, None),
    (32, dict(mal=17, msl=10, mrd=33, mqd=90, reg=50, aw=32, am=12, ar=5)),
    (33, dict(mal=8, msl=4, mrd=7, mqd=3, reg=10, aw=4, am=1, ar=1)),
    (35, dict(mal=2, msl=2, mrd=3, mqd=2, reg=4, aw=3, am=1, ar=1)),          # 16-bit bitmap: every position hits, record buffers overflow
    (36, dict(mal=1, msl=1, mrd=2, mqd=1, reg=2, aw=2, am=0, ar=1)),          # 4-bit bitmap
    (37, dict(mal=12, msl=7, mrd=40, mqd=40, reg=35, aw=15, am=7, ar=3)),     # direct 24-bit bitmap: probe scan, fold + exact check
    (38, dict(mal=11, msl=7, mrd=60, mqd=90, reg=35, aw=15, am=7, ar=3)),     # default k-mers, window too long for seed records
    (39, dict(mal=11, msl=7, mrd=10, mqd=57, reg=20, aw=15, am=7, ar=3)),     # seed records at their longest scan
    (40, dict(mal=11, msl=7, mrd=70, mqd=20, reg=35, aw=15, am=7, ar=3)),     # three-word rows from the start
    (34, dict(mal=11, msl=11, mrd=100, mqd=200, reg=35, aw=15, am=15, ar=3)),
    (35, dict(mal=15, msl=9, mrd=80, mqd=80, reg=100, aw=20, am=9, ar=4)),
    (41, dict(mal=15, msl=9, mrd=40, mqd=120, reg=60, aw=20, am=9, ar=4)),    # PrmLong with other run-time mrd / mqd / reg
])
def test_cuda_vs_oracle_on_fresh_families(seed, params):
    names, seqs = synth.families(3, 5, 4000, 9000, seed=seed, block=(200, 900))
    ss = hostio.SeqSet(names, seqs)
    got, _ = run_gpu(ss, params)
    want = Oracle().run_rows(ss, params)
    assert np.array_equal(got, want)


def test_sparse_csr_rows_and_batched_index_budget():
    ss = mg.get_set("example")
    n = len(ss)
    dense = GOLD["example_default"]
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        # sparse CSR through lzani_run (empty rows allowed)
        rng = np.random.default_rng(5)
        off = [0]
        q = []
        for r in range(n):
            cnt = int(rng.integers(0, 4))
            ids = rng.choice([x for x in range(n) if x != r], size=cnt, replace=False)
            q.extend(sorted(int(x) for x in ids))
            off.append(len(q))
        got = m.run(np.array(off, np.uint64), np.array(q, np.uint32))
        k = 0
        for r in range(n):
            for e in range(off[r], off[r + 1]):
                qq = q[e]
                assert got[k].tolist() == dense[r * (n - 1) + (qq if qq < r else qq - 1)].tolist()
                k += 1
        # a tiny index budget forces one reference per batch; results must not change
        m.set_index_budget(12 << 20)
        got2 = m.all2all()
        assert m.timing()["n_batches"] > 1
        assert np.array_equal(got2, dense)
        # row subset == the shard of one rank
        rows = lz.shard_rows(n, 1, 3)
        got3 = m.run_rows(rows)
        want3 = np.concatenate([dense[r * (n - 1):(r + 1) * (n - 1)] for r in rows])
        assert np.array_equal(got3, want3)
    finally:
        m.close()


def test_properties_at_bench_scale():
    """Size-independent properties on a slice of the bench workload (2k-genome config):
    identical genomes -> one region covering the whole genome; results independent of the
    row sharding; resident re-runs idempotent."""
    names, seqs = synth.viral(60, seed=synth.SEED)
    seqs = list(seqs) + [seqs[0].copy()]
    names = list(names) + ["dup_of_0"]
    ss = hostio.SeqSet(names, seqs)
    n = len(ss)
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        full = m.all2all()
        # identical pair (0, n-1): everything matches in one alignment unless the genome has N
        k = 0 * (n - 1) + (n - 2)
        L = int(ss.lens[0])
        if not (ss.codes[0] >= 4).any():
            assert full[k].tolist() == [L, 0, 1]
        # sharding invariance
        parts = {}
        for rank in range(4):
            rows = lz.shard_rows(n, rank, 4)
            res = m.run_rows(rows)
            for j, r in enumerate(rows):
                parts[int(r)] = res[j * (n - 1):(j + 1) * (n - 1)]
        again = np.concatenate([parts[r] for r in range(n)])
        assert np.array_equal(again, full)
        # resident repeat leaves the same results
        out = np.zeros_like(full)
        m.run_rows(np.arange(n, dtype=np.uint32), out=out, resident_repeat=2)
        assert np.array_equal(out, full)
        # checksum of checksums against the oracle on a bounded sample of rows
        rows = np.array([0, 7, n - 1], dtype=np.uint32)
        want = Oracle().run_rows(ss, None, rows=rows)
        got = m.run_rows(rows)
        assert np.array_equal(got, want)
    finally:
        m.close()


@pytest.mark.parametrize("seed", list(range(60000, 60018)))
def test_cuda_random_parameters_and_ragged_sets_differential(seed):
    """The generators of the CPU fuzz (tests/test_emul.py) through the CUDA library: default parameters, random windows
    with the default constants (lane walk, seed records or diagonal fallback), random everything (run-time parameter
    kernels); ragged low-complexity sets every other seed.  All pairs against the oracle."""
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
    got, _ = run_gpu(ss, params)
    want = Oracle().run_rows(ss, params, threads=4)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"{params}: {bad.size}/{len(want)} pairs differ, first {bad[:5]}"


def test_related_pairs_as_the_exception_and_as_the_rule(monkeypatch):
    """The lane walk hands related pairs to the warp-per-pair walk while they are at most a quarter of a chunk and keeps
    them otherwise (k_count_dense).  Both regimes, against the oracle on sampled rows and against the warp-per-pair
    walk (LZANI_LANE_WALK=0) on every pair: 160 genomes in families of 20 (12 % related pairs: handed over) and the
    first 40 of them in one chunk (49 % related: kept)."""
    names, seqs = synth.viral(160, seed=synth.SEED + 9)
    for count in (160, 40):
        ss = hostio.SeqSet(list(names[:count]), list(seqs[:count]))
        n = len(ss)
        monkeypatch.setenv("LZANI_LANE_WALK", "1")
        got, _ = run_gpu(ss, None)
        monkeypatch.setenv("LZANI_LANE_WALK", "0")
        warp, _ = run_gpu(ss, None)
        bad = np.nonzero((got != warp).any(axis=1))[0]
        assert bad.size == 0, f"{count} genomes: {bad.size} pairs differ between the two walks, first {bad[:5]}"
        rows = np.array([0, n // 3, n - 1], dtype=np.uint32)
        want = Oracle().run_rows(ss, None, rows=rows)
        assert np.array_equal(np.concatenate([got[r * (n - 1):(r + 1) * (n - 1)] for r in rows]), want), count


def test_sparse_filter_style_pair_list_config4():
    """BASELINE configs[3] shape at test size: a --flt-kmerdb-style sparse pair list (within-family
    pairs plus a few cross-family decoys, unsorted inside a row as filter rows are) through the CSR
    entry point, against the oracle."""
    names, seqs = synth.families(6, 6, 5000, 9000, seed=41, d_lo=0.02, d_hi=0.25, block=(200, 900))
    ss = hostio.SeqSet(names, seqs)
    n = len(ss)
    rng = np.random.default_rng(41)
    off, q = [0], []
    for r in range(n):
        fam = r // 6
        ids = [x for x in range(fam * 6, fam * 6 + 6) if x != r and rng.random() < 0.8]
        ids += [int(x) for x in rng.integers(0, n, size=2) if x != r and x // 6 != fam]
        rng.shuffle(ids)
        q.extend(ids)
        off.append(len(q))
    off, q = np.array(off, np.uint64), np.array(q, np.uint32)
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        got = m.run(off, q)
    finally:
        m.close()
    want = Oracle().run_rows(ss, None, row_off=off, q_ids=q)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("params", [mg.LONGP, None])
def test_megabase_genomes_config5(params):
    """BASELINE configs[4] shape at test size: bacterial-scale (1.2 Mb) related genomes with the
    long-genome parameter set (hashed bitmap, 64-bit keys, 2^23-slot tables) and with the defaults
    (dense hits, long matches, long extensions), against the oracle."""
    rng = np.random.default_rng(55)
    root = rng.integers(0, 4, size=1_200_000, dtype=np.uint8)
    seqs = [root, synth._mutate(rng, root, 0.04, block=(5000, 50000), p_block=1.0),
            synth._mutate(rng, root, 0.12, block=(5000, 50000), p_block=1.0)]
    ss = hostio.SeqSet(["root", "m4", "m12"], seqs)
    got, _ = run_gpu(ss, params)
    want = Oracle().run_rows(ss, params)
    assert np.array_equal(got, want), f"{got.tolist()} vs {want.tolist()}"


def test_five_megabase_genomes_full_config5_length():
    """BASELINE configs[4] at its real genome length: 5 Mb genomes, long-genome parameters -- a related pair (4 %
    divergence, block events), and an unrelated genome of the same size (so all bitmaps have one size and the hashed
    form of the all2all scan runs; 2^29-bit bitmaps, 2^25-slot tables) -- all six ordered pairs against the oracle."""
    rng = np.random.default_rng(56)
    root = rng.integers(0, 4, size=5_000_000, dtype=np.uint8)
    other = rng.integers(0, 4, size=4_900_000, dtype=np.uint8)
    seqs = [root, synth._mutate(rng, root, 0.04, block=(5000, 50000), p_block=1.0), other]
    ss = hostio.SeqSet(["root", "m4", "other"], seqs)
    got, t = run_gpu(ss, mg.LONGP)
    want = Oracle().run_rows(ss, mg.LONGP, threads=6)
    assert np.array_equal(got, want), f"{got.tolist()} vs {want.tolist()}"
    assert got[0, 0] > 4_000_000 and got[1, 0] < 100_000       # root-m4 aligns almost everywhere, root-other nowhere


@pytest.mark.parametrize("sname,params", [("example", None), ("example", mg.ODD), ("synth_small", mg.LONGP)])
def test_region_records_match_oracle(sname, params):
    """--out-alignment path (lzani_run_rows_regions, replaces CParser::get_parsing): counts and region
    records of every ordered pair against the oracle; a too small buffer is reported with the needed
    size and nothing written."""
    ss = mg.get_set(sname)
    n = len(ss)
    orc = Oracle()
    m = lz.Matcher(params=params, device=0)
    try:
        m.set_sequences(ss)
        out, off, regs = m.run_rows_regions(np.arange(n, dtype=np.uint32), cap=8)     # forces the retry path
        assert np.array_equal(out, m.all2all())
        k = 0
        total = 0
        for r in range(n):
            for q in range(n):
                if q == r:
                    continue
                want, st = orc.regions(ss, params, r, q)
                got = regs[int(off[k]):int(off[k + 1])]
                assert got.shape == want.shape and np.array_equal(got, want), (r, q)
                assert out[k].tolist() == st.tolist()
                total += len(want)
                k += 1
        assert total == len(regs) and total > 0
    finally:
        m.close()


@pytest.mark.parametrize("params", [mg.LONGP, dict(mal=17, msl=10, mrd=33, mqd=90, reg=50, aw=32, am=12, ar=5),
                                    dict(mal=12, msl=7, mrd=40, mqd=40, reg=35, aw=15, am=7, ar=3)])
def test_matrix_scan_over_hashed_bitmaps(params, monkeypatch):
    """Long anchors (2*mal > 24 bits): the bitmaps are hashed.  References of similar length get bitmaps of one size and
    the all2all scan then serves 32 of them per gather, like the direct form; a chunk whose bitmaps differ in size
    falls back to the probe scan.  Same integers either way, and as the oracle's."""
    names, seqs = synth.families(3, 6, 6000, 6100, seed=41, block=(200, 900))       # Dr within one power of two
    names2, seqs2 = synth.families(1, 4, 20000, 21000, seed=42)                     # a second bitmap size
    names2 = ["long_" + x for x in names2]
    hashed = 2 * params["mal"] > 24
    for label, (nm, sq) in dict(uniform=(list(names), list(seqs)), mixed=(list(names) + names2, list(seqs) + list(seqs2))).items():
        ss = hostio.SeqSet(nm, sq)
        monkeypatch.setenv("LZANI_MATRIX_SCAN", "1")
        got, t1 = run_gpu(ss, params)
        monkeypatch.setenv("LZANI_MATRIX_SCAN", "0")
        probe, t0 = run_gpu(ss, params)
        assert np.array_equal(got, probe), label
        assert np.array_equal(got, Oracle().run_rows(ss, params)), label
        if label == "uniform" or not hashed:
            assert t1["launches"] != t0["launches"], "the matrix form of the scan did not run"
        else:
            assert t1["launches"] == t0["launches"], "bitmaps of two sizes in one chunk must take the probe scan"


@pytest.mark.parametrize("env", [
    dict(LZANI_STAGED_WORDS="300"),                                  # many chunks, rows cut in the middle (probe scan for those)
    dict(LZANI_STAGED_WORDS="300", LZANI_STAGED_RECS="40"),          # records overflow: chunks halved, single pairs grow the buffers
    dict(LZANI_STAGED_WORDS="20000", LZANI_MATRIX_SCAN="0"),         # probe scan with the folded bitmap in shared memory
    dict(LZANI_STAGED="0"),                                          # one-kernel form
    dict(LZANI_LANE_WALK="0"),                                       # warp-per-pair walk for every pair (the lane walk hands it only the related ones)
    dict(LZANI_LANE_WALK="0", LZANI_STAGED_WORDS="300"),
])
def test_staged_chunking_paths(env, monkeypatch):
    """The host side of the staged parse under hostile capacities: chunk boundaries inside rows, record
    buffers that overflow (halving, growth), both scan kernels, dense and sparse lists -- same integers."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    for case in ("synth_small_default", "edge_default", "example_default"):
        if case not in mg.CASES:
            continue
        sname, params = mg.CASES[case]
        ss = mg.get_set(sname)
        got, t = run_gpu(ss, params)
        assert np.array_equal(got, GOLD[case]), (env, case)
    # sparse list (probe scan): every third pair of the example set
    sname, params = mg.CASES["example_default"]
    ss = mg.get_set(sname)
    n = len(ss)
    rows, off, q = [], [0], []
    for r in range(n):
        qs = [x for x in range(n) if x != r and (x + r) % 3 == 0]
        rows.append(r); q.extend(qs); off.append(len(q))
    m = lz.Matcher(params=params, device=0)
    try:
        m.set_sequences(ss)
        got = m.run_rows(np.array(rows, dtype=np.uint32), np.array(off, dtype=np.uint64), np.array(q, dtype=np.uint32))
    finally:
        m.close()
    full = GOLD["example_default"].reshape(n, n - 1, 3)
    want = np.array([full[r, x if x < r else x - 1] for r in range(n) for x in range(n) if x != r and (x + r) % 3 == 0])
    assert np.array_equal(got, want), env


def _two_devices():
    """Two device ids for the group tests: two GPUs when the box has them (peer access over NVLink), else the
    same GPU twice (two device contexts sharing it: same host logic)."""
    import torch
    return [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]


def test_group_context_owns_the_devices():
    """lzani_create(..., device_ids, n_devices > 1) -- ONE context, the library deals the reference rows
    over its devices (dense: r mod G; sparse: heaviest rows first), runs them from its own worker threads and
    hands back results in the caller's CSR order: dense all2all, sparse CSR with empty rows, row subsets, region
    records and the device-side out-filter (each device reads the other's rows in place) must equal the
    one-device answers."""
    ss = mg.get_set("example")
    n = len(ss)
    dense = GOLD["example_default"]
    single = lz.Matcher(device=0)
    grp = lz.Matcher(devices=_two_devices())
    try:
        single.set_sequences(ss)
        grp.set_sequences(ss)
        assert np.array_equal(grp.all2all(), dense)
        t = grp.timing()
        assert t["n_pairs"] == n * (n - 1) and t["launches"] > 0
        rng = np.random.default_rng(7)
        off, q = [0], []
        for r in range(n):
            cnt = int(rng.integers(0, 5))
            ids = rng.choice([x for x in range(n) if x != r], size=cnt, replace=False)
            q.extend(sorted(int(x) for x in ids))
            off.append(len(q))
        off, q = np.array(off, np.uint64), np.array(q, np.uint32)
        assert np.array_equal(grp.run(off, q), single.run(off, q))
        rows = np.array([5, 0, 9, 3, 7], dtype=np.uint32)
        assert np.array_equal(grp.run_rows(rows), single.run_rows(rows))
        # empty submit: legal, returns nothing
        assert grp.run_rows(np.zeros(0, np.uint32)).shape == (0, 3)
        assert single.run_rows(np.zeros(0, np.uint32)).shape == (0, 3)
        # regions
        a = single.run_rows_regions(np.arange(n, dtype=np.uint32))
        b = grp.run_rows_regions(np.arange(n, dtype=np.uint32), cap=8)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        # out-filter over results resident on both devices
        eff = ss.lens.astype(np.uint32)
        flt = dict(ani=0.9, qcov=0.3)
        out = np.zeros((n * (n - 1), 3), dtype=np.int32)
        single.run_rows(np.arange(n, dtype=np.uint32), out=out, resident_repeat=1)
        want = single.filter_pairs(eff, flt)
        grp.run_rows(np.arange(n, dtype=np.uint32), out=out, resident_repeat=1)
        got = grp.filter_pairs(eff, flt)
        key = lambda p: (int(p["a"]), int(p["b"]))
        assert sorted(map(key, got)) == sorted(map(key, want)) and len(got) > 0
        gd = {key(p): p for p in got}
        for p in want:
            g = gd[key(p)]
            assert g["ab"].tolist() == p["ab"].tolist() and g["ba"].tolist() == p["ba"].tolist() and g["lines"] == p["lines"]
    finally:
        single.close()
        grp.close()


def test_more_than_65535_sequences_dense_rows():
    """n > 65 535 sequences: the all2all matrix scan is off by construction (its grid's y dimension is the query
    id), dense rows go through the probe scan.  Tiny sequences, a few reference rows, against the oracle."""
    rng = np.random.default_rng(65)
    n = 66_000
    base = rng.integers(0, 4, size=90, dtype=np.uint8)
    seqs = []
    for i in range(n):
        s = base.copy() if i % 7 == 0 else rng.integers(0, 4, size=int(rng.integers(40, 120)), dtype=np.uint8)
        if i % 7 == 0:
            s[rng.integers(0, 90, size=3)] = rng.integers(0, 4, size=3, dtype=np.uint8)
        seqs.append(s)
    ss = hostio.SeqSet([f"s{i}" for i in range(n)], seqs)
    rows = np.array([0, 7, 65_999], dtype=np.uint32)
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        got = m.run_rows(rows)
    finally:
        m.close()
    want = Oracle().run_rows(ss, None, rows=rows)
    assert got.shape == (3 * (n - 1), 3)
    assert np.array_equal(got, want)
    assert int(got[:, 2].sum()) > 1000          # the near-copies of `base` do align
