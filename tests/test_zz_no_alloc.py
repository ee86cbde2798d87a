"""lzani_filter_dense_noalloc (include/lzani_b200.h; csrc/noalloc.cu): the allocation-free variant of the
--out-filter pre-filter.  No counterpart in the reference, so the
checker is the existing entry point, lzani_filter_pairs, on the same results -- bit-exact, like every other
integer result in this suite -- on two calls in a row (the second on a SMALLER set with other thresholds, so
anything left in the shared static scratch by the first call would show).

This file is named to be collected last.
"""
import ctypes as C
import importlib

import numpy as np
import pytest

import make_golden as mg

lz = importlib.import_module("lz-ani_b200")
KEYS = lz.FILTER_KEYS
MAX_SEQS = 2048          # LZANI_NOALLOC_MAX_SEQS

# (set, thresholds) of the two calls in a row
CALLS = [("vir61", dict(ani=0.9, qcov=0.3)), ("example", dict(tani=0.85, rcov=0.9, gani=0.8))]


def existing_entry_point(ss, thr):
    """(all2all results, surviving pairs) from lzani_run_rows_resident + lzani_filter_pairs."""
    n = len(ss)
    m = lz.Matcher(device=0)
    try:
        m.set_sequences(ss)
        res = m.all2all()
        m.run_rows(np.arange(n, dtype=np.uint32), None, None, resident_repeat=1)
        want = m.filter_pairs(ss.lens, thr)
    finally:
        m.close()
    return np.ascontiguousarray(res, dtype=np.int32), want


def call_noalloc(torch, res, eff_len, thr, cap, n=None, stream=None):
    """One call through the C ABI with caller-owned device buffers; returns (rc, message, n_out, pairs)."""
    lib = lz.load_library()
    n = len(eff_len) if n is None else n
    dev = torch.device("cuda", 0)
    d_res = torch.from_numpy(res.reshape(-1).view(np.int32).copy()).to(dev)
    d_len = torch.from_numpy(np.ascontiguousarray(eff_len, dtype=np.uint32).view(np.int32).copy()).to(dev)
    d_out = torch.zeros(max(cap, 1) * lz.PAIR_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    d_cnt = torch.full((1,), -7, dtype=torch.int64, device=dev)
    t = np.array([float(thr.get(k, 0.0)) for k in KEYS], dtype=np.float64)
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    rc = lib.lzani_filter_dense_noalloc(0, n, d_res.data_ptr(), d_len.data_ptr(), t.ctypes.data, d_out.data_ptr(), cap,
                                        d_cnt.data_ptr(), stream, err, len(err))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(0)[0]
    n_out = int(d_cnt.cpu().numpy().view(np.uint64)[0])
    pairs = d_out.cpu().numpy().view(lz.PAIR_DTYPE)[:cap]
    return rc, err.value.decode(), n_out, pairs, free0 - free1


@pytest.mark.gpu
def test_noalloc_filter_matches_existing_entry_point_two_calls_in_a_row():
    import torch
    assert lz.PAIR_DTYPE.itemsize == 36
    wants = [(mg.get_set(s), thr) + existing_entry_point(mg.get_set(s), thr) for s, thr in CALLS]
    assert len(wants[0][0]) > len(wants[1][0])          # the second call leaves stale scratch beyond its own size
    taken = []
    for ss, thr, res, want in wants:                    # the two calls in a row, nothing in between
        assert 0 < len(want) < len(ss) * (len(ss) - 1) // 2
        rc, msg, n_out, got, grew = call_noalloc(torch, res, ss.lens, thr, cap=len(want) + 5)
        assert rc == 0, msg
        assert n_out == len(want)
        for f in ("a", "b", "ab", "ba", "lines"):
            assert np.array_equal(got[f][:n_out], want[f]), f
        assert not got[n_out:].view(np.uint8).any()     # nothing written past the survivors
        taken.append(grew)
    assert taken[1] <= 0, f"device memory in use grew by {taken[1]} bytes across a call"
    # and once more on a side stream, the first set again
    ss, thr, res, want = wants[0]
    st = torch.cuda.Stream(device=0)
    rc, msg, n_out, got, _ = call_noalloc(torch, res, ss.lens, thr, cap=len(want), stream=st.cuda_stream)
    assert rc == 0 and n_out == len(want) and np.array_equal(got.tobytes(), want.tobytes())


@pytest.mark.gpu
def test_noalloc_filter_limits_cap_and_maximum_size():
    import torch
    from test_outfilter import numpy_masks
    ss, thr = mg.get_set("example"), CALLS[1][1]
    res, want = existing_entry_point(ss, thr)
    # cap below the survivor count: the full count is reported, the first cap pairs are written
    rc, msg, n_out, got, _ = call_noalloc(torch, res, ss.lens, thr, cap=2)
    assert rc == 0 and n_out == len(want) > 2
    assert np.array_equal(got.tobytes(), want[:2].tobytes())
    # beyond the maximum: a clear error, nothing launched
    rc, msg, n_out, _, _ = call_noalloc(torch, res, ss.lens, thr, cap=2, n=MAX_SEQS + 1)
    assert rc == 1 and "LZANI_NOALLOC_MAX_SEQS" in msg and str(MAX_SEQS + 1) in msg
    assert n_out == 2**64 - 7       # the count word was not touched
    # no pairs at all
    for n in (0, 1):
        rc, msg, n_out, _, _ = call_noalloc(torch, res, ss.lens, thr, cap=2, n=n)
        assert rc == 0 and n_out == 0
    # the maximum size (2048 sequences, 4 190 208 ordered pairs) on random integers, against the numpy float64
    # restatement of the decision function (tests/test_outfilter.py)
    n = MAX_SEQS
    rng = np.random.default_rng(5)
    big = np.stack([rng.integers(0, 60000, n * (n - 1)), rng.integers(0, 20000, n * (n - 1)),
                    rng.integers(0, 300, n * (n - 1))], 1).astype(np.int32)
    lens = rng.integers(30000, 60000, n).astype(np.uint32)
    a, b = np.triu_indices(n, 1)
    ab, ba = big[a * (n - 1) + (b - 1)], big[b * (n - 1) + a]
    masks = numpy_masks(dict(ani=0.74, qcov=0.9), ab, ba, lens[a], lens[b])
    keep = np.nonzero(masks)[0]
    assert 1000 < keep.size < a.size
    rc, msg, n_out, got, _ = call_noalloc(torch, big, lens, dict(ani=0.74, qcov=0.9), cap=keep.size)
    assert rc == 0, msg
    assert n_out == keep.size
    assert np.array_equal(got["a"], a[keep]) and np.array_equal(got["b"], b[keep])
    assert np.array_equal(got["lines"], masks[keep])
    assert np.array_equal(got["ab"], ab[keep]) and np.array_equal(got["ba"], ba[keep])
    # and the small case again right after the big one (the scratch is full of the big call's flags)
    rc, msg, n_out, got, _ = call_noalloc(torch, res, ss.lens, thr, cap=len(want))
    assert rc == 0 and n_out == len(want) and np.array_equal(got.tobytes(), want.tobytes())
