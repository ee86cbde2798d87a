#!/usr/bin/env python
"""BASELINE.json configs[3] and configs[4] AT SCALE, sharded over the GPUs of one box (one process per GPU, torchrun;
also runs as a plain single process).  One JSON line per config on rank 0.

  --config 4 : 100 000 synthetic phage genomes (5 000 families x 20, 30-70 kb), a --flt-kmerdb-style sparse pair list
               (within-family pairs whose simulated identity >= 0.7 plus 5 % random cross-family decoys).
               The edges go through lzani_build_pair_list (symmetric closure + id remap + CSR on the device, 8f-4); the
               rows are dealt over the ranks by lz.sharding.shard_rows_sparse (heaviest rows first, by query bases).
  --config 5 : 500 synthetic bacterial genomes (25 families x 20, 4.5-5.5 Mb, 1-15 % divergence, 5-50 kb block events),
               all2all with the long-genome parameters LONGP, rows r mod N.

Every rank checks a sample of ITS OWN pairs against the unmodified reference CParser (oracle/_ref) outside the timed
region.  Timing: device time of the passes (CUDA events in the library, max over ranks) and wall time of the public
call with host buffers (e2e).  Sets are generated once (rank 0, a pool of processes; lz.synth.fast_families) and cached
2-bit packed in /tmp for the other ranks.

    torchrun --nproc-per-node 8 tools/configs45.py --config 4
    python tools/configs45.py --config 5 --genomes 8           # one GPU, reduced
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
lz = importlib.import_module("lz-ani_b200")
hostio, synth = lz.hostio, lz.synth

LONGP = dict(mal=15, msl=9, mrd=80, mqd=80, reg=100, aw=20, am=9, ar=4)      # SURVEY 8d.5 long-genome params (recorded in BASELINE.md)


class PackedSet:
    """What Matcher.set_sequences and the checkers need of a sequence set (names, lengths, three_in_byte bytes) without
    keeping one byte per base of a 5 Gbp set around."""

    def __init__(self, names, lens, packed):
        self.names, self.lens, self.packed = list(names), np.asarray(lens, dtype=np.uint32), packed

    def __len__(self):
        return len(self.names)


def build_set(kind: str, n_genomes: int, rank: int, world: int, barrier, workers: int):
    """Returns (PackedSet in the reference's post-reorder order, divergences d, family ids, old->new id map)."""
    cache = f"/tmp/lzani_cfg_{kind}_{n_genomes}_{synth.SEED}.npz"
    if rank == 0 and not os.path.exists(cache):
        if kind == "phage":
            fam = (n_genomes + 19) // 20
            names, ds, seqs = synth.fast_families(fam, 20, 30_000, 70_000, seed=synth.SEED + 4, workers=workers)
        else:
            fam = (n_genomes + 19) // 20
            names, ds, seqs = synth.fast_families(fam, 20, 4_500_000, 5_500_000, d_lo=0.01, d_hi=0.15, seed=synth.SEED + 5,
                                                  block=(5_000, 50_000), workers=workers)
        names, ds, seqs = names[:n_genomes], ds[:n_genomes], seqs[:n_genomes]
        lens = np.array([s.size for s in seqs], dtype=np.int64)
        order = hostio.reorder(names, lens)              # new id i is old id order[i]
        packed = [hostio.pack3(seqs[i]) for i in order]
        del seqs
        plen = np.array([p.size for p in packed], dtype=np.int64)
        tmp = cache + f".{os.getpid()}.npz"
        np.savez(tmp, names=np.array([names[i] for i in order]), lens=lens[order], d=ds[order], order=order,
                 plen=plen, packed=np.concatenate(packed))
        os.replace(tmp, cache)
        del packed
    barrier()
    z = np.load(cache)
    plen = z["plen"]
    flat = z["packed"]
    offs = np.concatenate([[0], np.cumsum(plen)])
    packed = [flat[offs[i]:offs[i + 1]] for i in range(plen.size)]
    names = [str(x) for x in z["names"]]
    fam = np.array([int(x.split("_")[0][1:]) for x in names])
    order = z["order"]
    new_of_old = np.empty(order.size, dtype=np.uint32)
    new_of_old[order] = np.arange(order.size, dtype=np.uint32)
    return PackedSet(names, z["lens"], packed), z["d"], fam, new_of_old, order


def checker():
    from oracle_lib import Oracle, RefShim   # checker only
    return (RefShim(), "reference") if RefShim.available() else (Oracle(), "port")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True, choices=[4, 5])
    ap.add_argument("--genomes", type=int, default=0, help="default: 100000 (config 4) / 500 (config 5)")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--cpu-pairs", type=int, default=0, help="pairs of the parity / CPU sample per rank (default: 64 / 4)")
    ap.add_argument("--workers", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--pick", choices=["related", "unrelated"], default=None,
                    help="config 5, diagnostics: only the genomes of one family / one genome per family")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allred(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    workers = a.workers or max(1, (os.cpu_count() or 2) - 2)
    t_prep = time.perf_counter()
    if a.config == 4:
        G = a.genomes or 100_000
        ss, d, fam, new_of_old, order = build_set("phage", G, rank, world, barrier, workers)
        n = len(ss)
        # the kmer-db style edge list in the FILE's order (old ids, j < i): within-family pairs of simulated identity
        # >= 0.7 (1 - d_i - d_j: both members diverged independently from the family root) + 5 % cross-family decoys
        d_old = np.empty(n); d_old[order] = d
        fam_old = np.empty(n, dtype=np.int64); fam_old[order] = fam
        rng = np.random.default_rng(4)
        ei, ej = [], []
        for f0 in range(0, n, 20):                        # old ids of one family are consecutive
            ids = np.arange(f0, min(n, f0 + 20))
            ii, jj = np.nonzero(np.tril(np.ones((ids.size, ids.size), dtype=bool), -1))
            keep = 1.0 - d_old[ids[ii]] - d_old[ids[jj]] >= 0.7
            ei.append(ids[ii][keep]); ej.append(ids[jj][keep])
        ei, ej = np.concatenate(ei), np.concatenate(ej)
        n_decoy = max(1, ei.size // 20)
        di = rng.integers(1, n, n_decoy)
        dj = (rng.random(n_decoy) * di).astype(np.int64)
        keep = fam_old[di] != fam_old[dj]
        ei = np.concatenate([ei, di[keep]]).astype(np.uint32)
        ej = np.concatenate([ej, dj[keep]]).astype(np.uint32)
        params = None
    else:
        G = a.genomes or 500
        ss, d, fam, new_of_old, order = build_set("bact", G, rank, world, barrier, workers)
        if a.pick:
            idx = (np.nonzero(fam == fam[0])[0] if a.pick == "related"
                   else np.array([np.nonzero(fam == f)[0][0] for f in np.unique(fam)]))
            ss = PackedSet([ss.names[i] for i in idx], ss.lens[idx], [ss.packed[i] for i in idx])
        n = len(ss)
        params = LONGP
    prep_s = time.perf_counter() - t_prep

    m = lz.Matcher(params=params, device=local)
    t0 = time.perf_counter()
    m.set_sequences(ss)
    set_s = time.perf_counter() - t0
    pl_ms = 0.0
    if a.config == 4:
        off, q = m.build_pair_list(n, ei, ej, reorder_map=new_of_old)       # 8f-4: CSR on the device
        pl_ms = m.timing()["pairlist_ms"]
        shards = lz.sharding.shard_rows_sparse(off, ss.lens, q, world)
        rows = shards[rank]
        roff, rq = lz.sharding.sub_csr(rows, off, q)
        my_pairs = int(roff[-1])
        my_qb = int(ss.lens[rq].astype(np.int64).sum())
        run = lambda **kw: m.run_rows(rows, roff, rq, **kw)
    else:
        rows = lz.shard_rows(n, rank, world)
        my_pairs = rows.size * (n - 1)
        tl = int(ss.lens.astype(np.int64).sum())
        my_qb = int(sum(tl - int(ss.lens[r]) for r in rows))
        roff = rq = None
        run = lambda **kw: m.run_rows(rows, **kw)
    out = np.zeros((my_pairs, 3), dtype=np.int32)
    for _ in range(a.warmup):
        run(out=out, resident_repeat=1)
    barrier()
    dev_ms = parse_ms = index_ms = 0.0
    launches = batches = 0
    for _ in range(a.steps):
        run(out=out, resident_repeat=1)
        t = m.timing()
        dev_ms += t["total_ms"] - t["d2h_ms"]; parse_ms += t["parse_ms"]; index_ms += t["index_ms"]
        launches += t["launches"]; batches = t["n_batches"]
    barrier()
    t0 = time.perf_counter()
    if not a.no_e2e:
        run(out=out)                                      # e2e: pair list H2D, K2, K3, results D2H (sequences resident)
    barrier()
    e2e_s = time.perf_counter() - t0
    dev_ms_step = allred(dev_ms / a.steps, dist.ReduceOp.MAX if world > 1 else None)
    e2e_s = allred(e2e_s, dist.ReduceOp.MAX if world > 1 else None)
    total_pairs = allred(float(my_pairs), dist.ReduceOp.SUM if world > 1 else None)
    total_qb = allred(float(my_qb), dist.ReduceOp.SUM if world > 1 else None)

    # ---- parity on every rank: a sample of this rank's own pairs against the reference CParser
    chk, kind = checker()
    cores = os.cpu_count() or 1
    threads = max(1, cores // world)
    if a.config == 4:
        want_pairs = a.cpu_pairs or 64
        nz = np.nonzero(np.diff(roff.astype(np.int64)) > 0)[0]
        pick = nz[np.linspace(0, nz.size - 1, min(nz.size, max(1, want_pairs // 8))).astype(np.int64)] if nz.size else nz
        srows = rows[pick]
        soff = [0]; sq = []
        for j in pick:
            sq.append(rq[int(roff[j]):int(roff[j + 1])]); soff.append(soff[-1] + sq[-1].size)
        soff = np.array(soff, dtype=np.uint64); sq = np.concatenate(sq) if sq else np.zeros(0, np.uint32)
        got = np.concatenate([out[int(roff[j]):int(roff[j + 1])] for j in pick]) if len(pick) else np.zeros((0, 3), np.int32)
    else:
        want_pairs = a.cpu_pairs or 4
        nr = max(1, min(rows.size, threads, want_pairs))
        per = max(1, want_pairs // nr)
        pick = np.linspace(0, rows.size - 1, nr).astype(np.int64)
        srows = rows[pick]
        rng = np.random.default_rng(5 + rank)
        soff = [0]; sq = []; got = []
        for j in pick:
            r = int(rows[j])
            ids = np.sort(rng.choice(n - 1, size=min(per, n - 1), replace=False))
            got.append(out[j * (n - 1) + ids])
            sq.append((ids + (ids >= r)).astype(np.uint32)); soff.append(soff[-1] + ids.size)
        soff = np.array(soff, dtype=np.uint64); sq = np.concatenate(sq); got = np.concatenate(got)
    barrier()
    want = chk.run_rows(ss, params, rows=srows, row_off=soff, q_ids=sq, threads=threads)
    cpu_s = chk.last_seconds
    ok = bool(np.array_equal(got, want))
    if not ok:
        print(f"PARITY FAILURE on rank {rank}: {got[:4].tolist()} vs {want[:4].tolist()}", file=sys.stderr, flush=True)
    all_ok = allred(1.0 if ok else 0.0, dist.ReduceOp.MIN if world > 1 else None) > 0.5
    checked = allred(float(len(want)), dist.ReduceOp.SUM if world > 1 else None)
    cpu_s = allred(cpu_s, dist.ReduceOp.MAX if world > 1 else None)
    if rank == 0:
        line = {"config": "configs[3]: sparse phage pair list (--flt-kmerdb style, identity >= 0.7 + 5 % decoys)" if a.config == 4
                else "configs[4]: bacterial-scale all2all, long-genome parameters",
                "metric": "ordered_pairs_per_sec", "genomes": n, "n_gpus": world, "params": params or hostio.DEFAULT_PARAMS,
                "ordered_pairs": int(total_pairs), "steps": a.steps, "warmup": a.warmup,
                "value": total_pairs / (dev_ms_step / 1e3), "unit": "pairs/s", "ms_per_step": dev_ms_step,
                "query_gbp_per_sec": total_qb / (dev_ms_step / 1e3) / 1e9,
                "e2e": None if a.no_e2e else {"value": total_pairs / e2e_s, "unit": "pairs/s", "seconds": e2e_s},
                "rank0": {"index_ms": index_ms / a.steps, "parse_ms": parse_ms / a.steps, "launches_per_step": launches // a.steps,
                          "index_batches": batches, "pairs": my_pairs, "set_sequences_s": set_s, "pairlist_ms": pl_ms},
                "roofline": {"bound": "hbm", "achieved": (4.375 * total_qb + 12.0 * total_pairs) / (dev_ms_step / 1e3) / 1e9 / world,
                             "peak": json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0,
                             "unit": "GB/s per GPU"},
                "parity": {"ok": all_ok, "ranks": world, "pairs_checked_vs_reference_cpu": int(checked), "kind": kind},
                "cpu_baseline": {"value": checked / cpu_s, "unit": "pairs/s", "cores": threads * world, "kind": kind,
                                 "sample": f"{int(checked)} ordered pairs of the job, {threads} threads per rank, ranks side by side, {cpu_s:.1f} s"},
                "host_prep_s": prep_s}
        line["roofline"]["frac"] = line["roofline"]["achieved"] / line["roofline"]["peak"]
        print(json.dumps(line), flush=True)
    m.close()
    if world > 1:
        dist.destroy_process_group()
    if not all_ok:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
