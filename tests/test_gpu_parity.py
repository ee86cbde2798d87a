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
    (34, dict(mal=11, msl=11, mrd=100, mqd=200, reg=35, aw=15, am=15, ar=3)),
    (35, dict(mal=15, msl=9, mrd=80, mqd=80, reg=100, aw=20, am=9, ar=4)),
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
