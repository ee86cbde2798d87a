#!/usr/bin/env python
"""One-off generator of tests/golden/bench2000_rowsums.npz: the UNMODIFIED reference CParser
(oracle/_ref/libref_parser.so, built from /root/reference by oracle/Makefile) over ALL 2 000
reference rows of BASELINE config 2 (the bench.py set: 2 000 synthetic viral genomes, seed
20261104, default parameters, 3 998 000 ordered pairs).  Per reference row r it stores

    rowsums[r] = (sum nt_match, sum nt_mismatch, sum num_alns) over the 1 999 queries q != r
    rowhash[r] = sum_k (k+1) * (m_k + 3 x_k + 7 n_k)  mod 2^64, k = position of the pair in the row

so that every bench.py run can check all 3 998 000 CUDA results against the reference, not a sample.
About 40 CPU-minutes on 8 cores; resumable (progress in /tmp).

    python tests/golden/make_bench_rowsums.py [--genomes 2000] [--threads 6]
"""
import argparse
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def row_digest(res: np.ndarray, n_rows: int):
    """res: (n_rows * per_row, 3) int32 in CSR order -> (rowsums int64 [n_rows,3], rowhash uint64 [n_rows])."""
    r = res.reshape(n_rows, -1, 3).astype(np.int64)
    sums = r.sum(axis=1)
    k = np.arange(1, r.shape[1] + 1, dtype=np.uint64)[None, :]
    v = (r[:, :, 0] + 3 * r[:, :, 1] + 7 * r[:, :, 2]).astype(np.uint64)
    with np.errstate(over="ignore"):
        h = (k * v).sum(axis=1, dtype=np.uint64)
    return sums, h


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genomes", type=int, default=2000)
    ap.add_argument("--threads", type=int, default=6)
    ap.add_argument("--block", type=int, default=48, help="reference rows per checkpoint")
    args = ap.parse_args()
    import bench
    from oracle_lib import RefShim
    ss = bench.make_set(args.genomes)
    n = len(ss)
    chk = RefShim()
    tmp = f"/tmp/bench{n}_rowsums_partial.npz"
    sums = np.zeros((n, 3), dtype=np.int64)
    hsh = np.zeros(n, dtype=np.uint64)
    done = 0
    if os.path.exists(tmp):
        z = np.load(tmp)
        sums, hsh, done = z["rowsums"], z["rowhash"], int(z["done"])
    while done < n:
        rows = np.arange(done, min(n, done + args.block), dtype=np.uint32)
        res = chk.run_rows(ss, None, rows=rows, threads=args.threads)
        s, h = row_digest(res, rows.size)
        sums[rows] = s
        hsh[rows] = h
        done = int(rows[-1]) + 1
        np.savez(tmp, rowsums=sums, rowhash=hsh, done=done)
        print(f"{done}/{n} rows, {chk.last_seconds:.1f} s for the last {rows.size}", flush=True)
    out = os.path.join(HERE, f"bench{n}_rowsums.npz")
    np.savez_compressed(out, rowsums=sums, rowhash=hsh, genomes=n, seed=bench.synth.SEED,
                        lens=ss.lens.astype(np.uint32))
    print("wrote", out, "totals", sums.sum(axis=0).tolist())


if __name__ == "__main__":
    main()
