"""world_size-2 `gloo` test of the multi-GPU host logic: rows dealt by shard_rows /
shard_rows_sparse, per-rank results gathered and merged into CSR order.  No GPU here, so each rank's
"device" is the oracle (TEST stand-in for lzani_run_rows; the product path has no CPU fallback)."""
import importlib
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle_lib import Oracle, hostio
import make_golden as mg

lz = importlib.import_module("lz-ani_b200")
sharding = lz.sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, sparse, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ss = mg.get_set("example")
        n = len(ss)
        orc = Oracle()
        if not sparse:
            rows = sharding.shard_rows(n, rank, world)
            local = orc.run_rows(ss, None, rows=rows, threads=2)
            merged = sharding.gather_results(rows, local, n, n - 1)
        else:
            rng = np.random.default_rng(3)
            off = [0]
            qs = []
            for r in range(n):
                ids = rng.choice([x for x in range(n) if x != r], size=int(rng.integers(0, 6)), replace=False)
                qs.extend(int(x) for x in ids)
                off.append(len(qs))
            off, qs = np.array(off, np.uint64), np.array(qs, np.uint32)
            parts = sharding.shard_rows_sparse(off, ss.lens, qs, world)
            rows = parts[rank]
            soff, sq = sharding.sub_csr(rows, off, qs)
            local = orc.run_rows(ss, None, rows=rows, row_off=soff, q_ids=sq, threads=2)
            merged = sharding.gather_results(rows, local, n, np.diff(off.astype(np.int64)))
            if rank == 0:
                want = orc.run_rows(ss, None, row_off=off, q_ids=qs, threads=2)
                q.put(bool(np.array_equal(merged, want)) and sorted(np.concatenate(parts).tolist()) == list(range(n)))
        if not sparse and rank == 0:
            want = np.load(os.path.join(os.path.dirname(mg.__file__), "golden.npz"))["example_default"]
            q.put(bool(np.array_equal(merged, want)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _run(sparse):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, sparse, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def test_dense_rows_two_ranks_gloo():
    _run(False)


def test_sparse_rows_two_ranks_gloo():
    _run(True)


def test_merge_rows_detects_missing_and_duplicate_rows():
    import pytest
    res = np.zeros((2, 3), np.int32)
    with pytest.raises(AssertionError):
        sharding.merge_rows([(np.array([0]), res[:1])], 2, 1)
    with pytest.raises(AssertionError):
        sharding.merge_rows([(np.array([0]), res[:1]), (np.array([0]), res[:1])], 2, 1)
    assert sharding.merge_rows([(np.array([1]), res[:1] + 5), (np.array([0]), res[:1])], 2, 1)[1, 0] == 5
