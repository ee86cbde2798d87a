"""Host-side sharding of the pair list over ranks: by REFERENCE ROW, the reference's own
task unit (lz_matcher.cpp:200-209).  Sequences are replicated on every GPU; each rank parses the
rows it owns; per-rank results are concatenated on the host.  There is no collective on the data
path -- `gather_results` is the only communication and it moves the 12-byte results only.
"""
from __future__ import annotations

import numpy as np


def shard_rows(n_refs: int, rank: int, world: int) -> np.ndarray:
    """Reference rows of one rank: r mod world == rank over the length-descending ids, so every
    rank gets the same length mix."""
    return np.arange(rank, n_refs, world, dtype=np.uint32)


def shard_rows_sparse(row_off: np.ndarray, lens: np.ndarray, q_ids: np.ndarray, world: int):
    """Sparse pair lists (--flt-kmerdb): greedy assignment of rows to the least loaded rank by
    sum of query bases, heaviest rows first.  Returns a list of row-id arrays (ascending)."""
    row_off = np.asarray(row_off, dtype=np.int64)
    n = row_off.size - 1
    cs = np.concatenate([[0], np.cumsum(np.asarray(lens, dtype=np.int64)[np.asarray(q_ids, dtype=np.int64)])])
    cost = cs[row_off[1:]] - cs[row_off[:-1]]
    load = np.zeros(world, dtype=np.int64)
    owner = np.zeros(n, dtype=np.int64)
    for r in np.argsort(-cost, kind="stable"):
        g = int(np.argmin(load))
        owner[r] = g
        load[g] += cost[r]
    return [np.nonzero(owner == g)[0].astype(np.uint32) for g in range(world)]


def sub_csr(rows: np.ndarray, row_off: np.ndarray, q_ids: np.ndarray):
    """CSR restricted to `rows` (in that order)."""
    row_off = np.asarray(row_off, dtype=np.uint64)
    off = np.zeros(len(rows) + 1, dtype=np.uint64)
    parts = []
    for j, r in enumerate(rows):
        a, b = int(row_off[r]), int(row_off[r + 1])
        parts.append(np.asarray(q_ids[a:b], dtype=np.uint32))
        off[j + 1] = off[j] + np.uint64(b - a)
    q = np.concatenate(parts) if parts else np.zeros(0, np.uint32)
    return off, q


def merge_rows(pieces, n_refs: int, row_len) -> np.ndarray:
    """pieces: iterable of (rows, results) per rank, results (sum(row_len[rows]), 3) in the rank's row
    order.  Returns the (sum(row_len), 3) array in global reference-row order (the CSR order of
    lzani_run).  row_len: int (dense all2all: n-1) or array of per-row pair counts."""
    rl = np.full(n_refs, row_len, dtype=np.int64) if np.isscalar(row_len) else np.asarray(row_len, dtype=np.int64)
    off = np.concatenate([[0], np.cumsum(rl)])
    out = np.zeros((int(off[-1]), 3), dtype=np.int32)
    seen = np.zeros(n_refs, dtype=bool)
    for rows, res in pieces:
        k = 0
        for r in np.asarray(rows, dtype=np.int64):
            assert not seen[r], f"row {r} owned by two ranks"
            seen[r] = True
            out[off[r]:off[r + 1]] = res[k:k + rl[r]]
            k += int(rl[r])
        assert k == len(res)
    assert seen.all() or n_refs == 0, "rows missing from the shards"
    return out


def gather_results(rows: np.ndarray, local: np.ndarray, n_refs: int, row_len, dst: int = 0):
    """torch.distributed gather of the per-rank results to `dst`; returns the merged array there,
    None elsewhere.  Works with any backend (gloo in the CPU tests, nccl under torchrun)."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    objs = [None] * world if rank == dst else None
    dist.gather_object((np.asarray(rows), np.asarray(local)), objs, dst=dst)
    if rank != dst:
        return None
    return merge_rows(objs, n_refs, row_len)
