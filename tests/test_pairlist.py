"""Sparse pair list built on the device (reference CFilter::load_filter symmetric closure,
src/filter.cpp:243-280, and CFilter::reorder_items, :293-345).

`restate_filter` below is a plain-Python restatement of what the reference leaves in CFilter::filter
after load + reorder.  It is pinned on the CPU against the reference CLI: the (ridx, qidx) lines of the
golden TSV the unmodified CLI wrote with `--flt-kmerdb example/fltr.txt 0.9` are exactly its entries.
The GPU tests then check lzani_build_pair_list against it (row multisets, ascending in-row order) and
that parsing with the device-built list reproduces the reference's integers for those pairs.
"""
import gzip
import importlib
import os

import numpy as np
import pytest

import make_golden as mg

lz = importlib.import_module("lz-ani_b200")
GOLD_DIR = os.path.dirname(mg.__file__)
GOLD = np.load(os.path.join(GOLD_DIR, "golden.npz"))
FLTR = os.path.join(GOLD_DIR, "example_fltr.txt")


def parse_kmerdb(path, thr):
    """names (file order) and the edges (i, j) whose value passes `thr` (filter.cpp:34-42, 64-84):
    header `<x>,name1,name2,...`; row i = `name_i,<j+1>:<val>,...`."""
    lines = open(path).read().split("\n")
    names = [x for x in lines[0].split(",")[1:]]
    ei, ej = [], []
    i = 0
    for line in lines[1:]:
        if len(line) <= 2:
            continue
        for part in line.split(",")[1:]:
            e = part.split(":")
            if len(e) == 2 and float(e[1]) >= thr:
                ei.append(i)
                ej.append(int(e[0]) - 1)
        i += 1
    return names, np.array(ei, np.uint32), np.array(ej, np.uint32)


def restate_filter(n, ei, ej, reorder_map=None):
    """rows[new id] = sorted list of new ids: every edge in both rows, then the id permutation."""
    mp = np.arange(n) if reorder_map is None else np.asarray(reorder_map)
    rows = [[] for _ in range(n)]
    for i, j in zip(ei.tolist(), ej.tolist()):
        rows[mp[i]].append(int(mp[j]))
        rows[mp[j]].append(int(mp[i]))
    return [sorted(r) for r in rows]


def example_setup():
    ss = mg.get_set("example")                  # post-reorder ids
    names, ei, ej = parse_kmerdb(FLTR, 0.9)
    names = [x for x in names if x]
    reorder_map = np.array([ss.names.index(x) for x in names], np.uint32)   # file order -> post-reorder id
    return ss, ei, ej, reorder_map


def test_restatement_is_what_the_reference_cli_compares():
    ss, ei, ej, mp = example_setup()
    rows = restate_filter(len(ss), ei, ej, mp)
    mine = {(r, q) for r in range(len(ss)) for q in rows[r]}
    txt = gzip.open(os.path.join(GOLD_DIR, "example_filter09.ani.tsv.gz")).read().decode().splitlines()
    ref = {(int(l.split("\t")[1]), int(l.split("\t")[0])) for l in txt[1:]}
    assert mine == ref and len(mine) == 2 * ei.size


@pytest.mark.gpu
def test_device_pair_list_equals_restatement_and_parses_to_golden():
    ss, ei, ej, mp = example_setup()
    n = len(ss)
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        off, q = m.build_pair_list(n, ei, ej, mp)
        rows = restate_filter(n, ei, ej, mp)
        assert off[0] == 0 and off[-1] == 2 * ei.size
        for r in range(n):
            assert q[int(off[r]):int(off[r + 1])].tolist() == rows[r]
        got = m.run(off, q)
        assert m.timing()["pairlist_ms"] >= 0
    finally:
        m.close()
    dense = GOLD["example_default"]
    k = 0
    for r in range(n):
        for qq in rows[r]:
            assert got[k].tolist() == dense[r * (n - 1) + (qq if qq < r else qq - 1)].tolist()
            k += 1


@pytest.mark.gpu
@pytest.mark.parametrize("n,n_edges,with_map", [(1, 0, False), (5, 0, True), (1000, 20000, True), (70000, 300000, True),
                                                 (300, 5000, False)])
def test_device_pair_list_random(n, n_edges, with_map):
    rng = np.random.default_rng(n + n_edges)
    ei = rng.integers(0, n, n_edges).astype(np.uint32)
    ej = rng.integers(0, n, n_edges).astype(np.uint32)      # duplicates and self edges on purpose: nothing is dropped
    mp = rng.permutation(n).astype(np.uint32) if with_map else None
    m = lz.Matcher(device=0)
    try:
        off, q = m.build_pair_list(n, ei, ej, mp)
    finally:
        m.close()
    a = ei.astype(np.int64) if mp is None else mp[ei].astype(np.int64)
    b = ej.astype(np.int64) if mp is None else mp[ej].astype(np.int64)
    r = np.concatenate([a, b])
    c = np.concatenate([b, a])
    order = np.lexsort((c, r))
    assert np.array_equal(q, c[order].astype(np.uint32))
    want_off = np.concatenate([[0], np.cumsum(np.bincount(r, minlength=n))]).astype(np.uint64)
    assert np.array_equal(off, want_off)


@pytest.mark.gpu
def test_device_pair_list_rejects_bad_ids():
    m = lz.Matcher(device=0)
    try:
        with pytest.raises(lz.LzaniError, match="sequence id"):
            m.build_pair_list(10, np.array([3, 11], np.uint32), np.array([1, 2], np.uint32))
    finally:
        m.close()
