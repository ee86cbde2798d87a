#!/usr/bin/env python
"""bench.py -- the LZ-ANI all2all pair-comparison hot path on B200 (contract: see DESIGN.md).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--genomes G] [--scaling weak|strong] [--config 2|3]

Workload (BASELINE.json configs[1]): seeded synthetic viral genomes, ~50 kb, families of 20 with 1-30 %
divergence + indels + block events (lz-ani_b200/synth.py, seed 20261104), all2all, default parameters.  One STEP
= one pass of the hot path over the whole pair list of this rank: reference index build (K2) + the staged parse
(K3: per chunk of pairs the data-parallel scan / expand / probe / spec kernels and the walk: one lane per pair,
k_walk_lanes; related pairs one warp per pair, k_walk_pairs).  N=1: G = 2000 genomes -> 3 998 000 ordered pairs
per step.  N>1 (one process per GPU, torchrun):
  --scaling weak (default): the set grows to G_N = round(G*sqrt(N)) genomes so that every rank keeps ~G*(G-1) pairs;
  --scaling strong / --config 3: a FIXED set (config 3 = BASELINE configs[2]: 20 000 genomes, 3.9998e8 ordered pairs)
    whatever N is.
Rank r owns reference rows r, r+N, ... (pair list sharded by reference block, sequences replicated, no collective
on the data path).

  value : ordered pairs/s, whole job, sequences + pair list resident in HBM, device-timed (CUDA events on the
          library's stream), max over ranks.
  e2e   : the same metric through the public C ABI with HOST buffers each step: lzani_set_sequences (H2D + K1
          pack) + lzani_run_rows (pair list H2D, K2, K3, results D2H), wall-clock around the calls, max over ranks.
  parity: EVERY rank, at every N, compares a sample of its own rows with the UNMODIFIED reference CParser
          (oracle/_ref) outside the timed region; at 2 000 genomes every result of every rank is also checked against
          the reference's committed per-row digests (tests/golden/bench2000_rowsums.npz: all 3 998 000 pairs).
  cpu_baseline / --impl reference : the reference CParser on the box's host cores, bounded sample.
"""
from __future__ import annotations

import argparse
import hashlib
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
lz = importlib.import_module("lz-ani_b200")
hostio, synth = lz.hostio, lz.synth

METRIC = "all2all_ordered_pairs_per_sec"
UNIT = "pairs/s"
BYTES_PER_QUERY_BASE = 4.375      # 3 bits of sequence + one 4-byte anchor slot per query position
BYTES_PER_REF_POS = 8.375         # index build: 3 bits read + 4 B + 4 B written per R-form position
REF_BUILD = "g++ -O3 (oracle/Makefile; the reference's own makefile adds -march=native and links mimalloc: not used, the binary travels to another CPU)"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def csrc_digest() -> str:
    """Digest of the kernel sources the library was built from: ties profiles/traffic.json to a build."""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "lz-ani_b200", "csrc")
    for fn in sorted(os.listdir(d)):
        if fn.endswith((".cu", ".h")):
            h.update(fn.encode())
            h.update(open(os.path.join(d, fn), "rb").read())
    return h.hexdigest()[:16]


def measured_traffic(n_genomes: int, world: int):
    """DRAM bytes of the parse stage per step from the committed ncu capture (profiles/traffic.json, written by
    tools/traffic.py from `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`) -- only when that capture was
    taken from THESE kernel sources on THIS workload; otherwise null."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        t = json.load(open(p))
        if int(t["genomes"]) == n_genomes and world == 1 and t.get("csrc_sha1") == csrc_digest():
            return float(t["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


def _pack_set(names, seqs):
    """2 bits per base + the positions of the N symbols: a 20 000-genome set is 250 MB instead of 1 GB."""
    flat = np.concatenate(seqs)
    npos = np.nonzero(flat > 3)[0].astype(np.int64)
    c = (flat & 3).astype(np.uint8)
    c[npos] = 0
    pad = (-c.size) % 4
    c = np.concatenate([c, np.zeros(pad, np.uint8)]).reshape(-1, 4)
    packed = (c[:, 0] | (c[:, 1] << 2) | (c[:, 2] << 4) | (c[:, 3] << 6)).astype(np.uint8)
    return dict(names=np.array(names), lens=np.array([s.size for s in seqs], dtype=np.int64), packed=packed, npos=npos,
                nval=flat[npos])


def _unpack_set(z):
    lens = z["lens"]
    p = z["packed"]
    flat = np.empty((p.size, 4), dtype=np.uint8)
    for k in range(4):
        flat[:, k] = (p >> (2 * k)) & 3
    flat = flat.reshape(-1)[: int(lens.sum())]
    flat[z["npos"]] = z["nval"]
    offs = np.concatenate([[0], np.cumsum(lens)])
    return [str(x) for x in z["names"]], [flat[offs[i]:offs[i + 1]] for i in range(len(lens))]


def make_set(n_genomes: int):
    """The seeded synthetic set (synth.viral), cached: /tmp on this box, or .synth_cache/ in the repository tree (a
    pre-generated copy travels with the snapshot: generating 20 000 genomes takes ~100 s of one core)."""
    fn = f"lzani_synth_{synth.SEED}_{n_genomes}.v2.npz"
    names = seqs = None
    for cache in (os.path.join("/tmp", fn), os.path.join(ROOT, ".synth_cache", fn)):
        if os.path.exists(cache):
            try:
                names, seqs = _unpack_set(np.load(cache))
                break
            except Exception:
                names = seqs = None
    if names is None:
        names, seqs = synth.viral(n_genomes)
        try:
            cache = os.path.join("/tmp", fn)
            tmp = cache + f".{os.getpid()}.npz"
            np.savez(tmp, **_pack_set(names, seqs))
            os.replace(tmp, cache)
        except Exception:
            pass
    # ids in the reference's post-reorder order (length descending, name ascending)
    order = hostio.reorder(names, [s.size for s in seqs])
    return hostio.SeqSet([names[i] for i in order], [seqs[i] for i in order])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def n_genomes_for(args, world: int) -> int:
    if args.config == 3:
        return 20000
    if args.scaling == "strong":
        return args.genomes
    return int(round(args.genomes * math.sqrt(world)))


def sample_csr(n: int, rows: np.ndarray, per_row: int, seed: int):
    """A bounded sample of the all2all pair list: for each of `rows`, `per_row` query ids (ascending, != row),
    all of them when per_row >= n - 1.  Returns (row_off, q_ids)."""
    rng = np.random.default_rng(seed)
    off, q = [0], []
    for r in rows:
        r = int(r)
        if per_row >= n - 1:
            ids = np.concatenate([np.arange(0, r), np.arange(r + 1, n)])
        else:
            ids = np.sort(rng.choice(n - 1, size=per_row, replace=False))
            ids = ids + (ids >= r)
        q.append(ids.astype(np.uint32))
        off.append(off[-1] + ids.size)
    return np.array(off, dtype=np.uint64), (np.concatenate(q) if q else np.zeros(0, np.uint32))


def checker():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle, RefShim   # checker / baseline only, never on the product path
    return (RefShim(), "reference") if RefShim.available() else (Oracle(), "port")


def reference_arm(args, rank, world):
    """The reference's own CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    ss = make_set(n_genomes_for(args, args.gpus))
    n = len(ss)
    chk, kind = checker()
    # a bounded sample per step: one reference row per host thread, ~cpu_pairs ordered pairs in all
    n_rows = max(1, min(n, args.cpu_rows or cores))
    per_row = max(1, min(n - 1, args.cpu_pairs // n_rows))
    pairs = n_rows * per_row
    rows_all = np.linspace(0, n - 1, n_rows * (args.warmup + args.steps)).astype(np.uint32)
    times = []
    for s in range(args.warmup + args.steps):
        rows = rows_all[s * n_rows:(s + 1) * n_rows]
        off, q = sample_csr(n, rows, per_row, seed=1000 + s)
        chk.run_rows(ss, None, rows=rows, row_off=off, q_ids=q, threads=cores)
        if s >= args.warmup:
            times.append(chk.last_seconds)
    ms = 1e3 * float(np.mean(times))
    v = pairs / (ms / 1e3)
    qb = float(ss.lens.sum()) / n * pairs
    sample = f"{n_rows} reference rows x {per_row} queries per step on {cores} host threads; built with {REF_BUILD}"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"synthetic viral all2all, {n} genomes ~50 kb, default params; each step = {n_rows} reference rows x {per_row} queries ({pairs} ordered pairs) on {cores} host threads",
                       "genomes": n, "params": hostio.DEFAULT_PARAMS},
            "query_gbp_per_sec": qb / (ms / 1e3) / 1e9,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def golden_rowsums(n: int, ss):
    """The reference's digests of every reference row of the 2 000-genome set, if this IS that set."""
    p = os.path.join(ROOT, "tests", "golden", f"bench{n}_rowsums.npz")
    if not os.path.exists(p):
        return None
    z = np.load(p)
    if int(z["genomes"]) != n or int(z["seed"]) != synth.SEED or not np.array_equal(z["lens"], ss.lens.astype(np.uint32)):
        return None
    return z["rowsums"], z["rowhash"]


def row_digest(res: np.ndarray, n_rows: int):
    """(n_rows * per_row, 3) int32 in CSR order -> per-row sums and position-weighted hashes (the same digests
    tests/golden/make_bench_rowsums.py took of the reference's results)."""
    r = res.reshape(n_rows, -1, 3).astype(np.int64)
    sums = r.sum(axis=1)
    k = np.arange(1, r.shape[1] + 1, dtype=np.uint64)[None, :]
    v = (r[:, :, 0] + 3 * r[:, :, 1] + 7 * r[:, :, 2]).astype(np.uint64)
    with np.errstate(over="ignore"):
        h = (k * v).sum(axis=1, dtype=np.uint64)
    return sums, h


def peer_group_check(ss_small):
    """One process, ONE library context driving devices 0 and 1 (lzani_create with n_devices = 2): rows dealt by the
    library, and the device-side out-filter in which each GPU reads the other's rows in place over NVLink peer
    access -- against the same calls on one device."""
    n = len(ss_small)
    rows = np.arange(n, dtype=np.uint32)
    one, grp = lz.Matcher(device=0), lz.Matcher(devices=[0, 1])
    try:
        one.set_sequences(ss_small)
        grp.set_sequences(ss_small)
        a = one.all2all()
        ok = bool(np.array_equal(a, grp.all2all()))
        eff = ss_small.lens.astype(np.uint32)
        flt = dict(ani=0.8, qcov=0.2)
        out = np.zeros_like(a)
        one.run_rows(rows, out=out, resident_repeat=1)
        want = one.filter_pairs(eff, flt)
        grp.run_rows(rows, out=out, resident_repeat=1)
        got = grp.filter_pairs(eff, flt)
        key = lambda p: (int(p["a"]), int(p["b"]), p["ab"].tolist(), p["ba"].tolist(), int(p["lines"]))
        ok = ok and sorted(map(key, got)) == sorted(map(key, want)) and len(want) > 0
        return ok, len(want)
    finally:
        one.close()
        grp.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--genomes", type=int, default=2000, help="genomes at N=1 (BASELINE configs[1]: 2000)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="weak: G*sqrt(N) genomes; strong: G genomes at every N")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3], help="3 = BASELINE configs[2]: 20 000 genomes, fixed job (implies --scaling strong)")
    ap.add_argument("--cpu-rows", type=int, default=0, help="reference rows per CPU-baseline step (default: host cores)")
    ap.add_argument("--cpu-pairs", type=int, default=32000, help="ordered pairs per CPU-baseline / parity sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.config == 3:
        args.scaling = "strong"
    rank, world, local = dist_env()
    if world != args.gpus and world != 1:
        args.gpus = world
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allreduce(x: float, op) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    allmax = lambda x: allreduce(x, dist.ReduceOp.MAX) if world > 1 else x
    allsum = lambda x: allreduce(x, dist.ReduceOp.SUM) if world > 1 else x
    allmin = lambda x: allreduce(x, dist.ReduceOp.MIN) if world > 1 else x

    G = n_genomes_for(args, world)
    if world > 1 and rank != 0:
        barrier()                     # rank 0 generates (and caches) the set first
    ss = make_set(G)
    if world > 1 and rank == 0:
        barrier()
    n = len(ss)
    rows = lz.shard_rows(n, rank, world)
    my_pairs = rows.size * (n - 1)
    total_len = int(ss.lens.sum())
    my_qbases = int(sum(total_len - int(ss.lens[r]) for r in rows))

    m = lz.Matcher(device=local)
    m.set_sequences(ss)
    out = np.zeros((my_pairs, 3), dtype=np.int32)

    # ---- value: resident, device-timed
    for _ in range(args.warmup):
        m.run_rows(rows, resident_repeat=1, out=out)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    dev_ms = parse_ms = index_ms = 0.0
    launches = 0
    r_bases = 0
    for _ in range(args.steps):
        m.run_rows(rows, resident_repeat=1, out=out)
        t = m.timing()
        dev_ms += t["total_ms"] - t["d2h_ms"]
        parse_ms += t["parse_ms"]
        index_ms += t["index_ms"]
        launches += t["launches"]
        r_bases = t["r_bases"]
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop()
    dev_ms_step = allmax(dev_ms / args.steps)
    wall_ms_step = allmax(wall_ms / args.steps)
    parse_ms_step = allmax(parse_ms / args.steps)
    total_pairs = allsum(float(my_pairs))
    total_qbases = allsum(float(my_qbases))
    value = total_pairs / (dev_ms_step / 1e3)
    checksum = out.astype(np.int64).sum(axis=0)
    resident = out.copy() if out.nbytes <= (1 << 30) else None

    # ---- e2e: host buffers through the C ABI every step
    e2e = None
    if not args.no_e2e:
        h2d = d2h = 0
        for _ in range(1):
            m.set_sequences(ss)
            m.run_rows(rows, out=out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            m.set_sequences(ss)
            h2d = m.timing()["h2d_bytes"]
            m.run_rows(rows, out=out)
            t = m.timing()
            h2d += t["h2d_bytes"]
            d2h = t["d2h_bytes"]
            launches += t["launches"] + 1
        barrier()
        e2e_ms_step = allmax(1e3 * (time.perf_counter() - t0) / args.steps)
        e2e = {"value": total_pairs / (e2e_ms_step / 1e3), "unit": UNIT, "ms_per_step": e2e_ms_step,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}
        assert out.astype(np.int64).sum(axis=0).tolist() == checksum.tolist(), "e2e results differ from the resident run"
        if resident is not None:
            assert np.array_equal(out, resident), "e2e results differ from the resident run"

    # ---- roofline of the parse stage, per step.  The parse is a pipeline of kernels per chunk of pairs (scan, expand,
    # probe, spec, walk); SURVEY 8d's algorithmic bytes (4.375 B per query base + 12 B per pair) are those of the WHOLE
    # parse of a pair, so they are divided by the time of the whole stage (CUDA events around it on the library's
    # stream), not by one kernel's share of it.
    peak, peak_src = load_peaks()
    alg_bytes = BYTES_PER_QUERY_BASE * my_qbases + 12.0 * my_pairs
    achieved = alg_bytes / (parse_ms / args.steps / 1e3) / 1e9 if parse_ms > 0 else 0.0
    staged = os.environ.get("LZANI_STAGED", "1") != "0"
    roofline = {"bound": "hbm", "kernel": "parse stage: k_scan_matrix + k_expand_hits + k_hit_probe + k_hit_spec (largest) + k_walk_lanes + k_walk_pairs" if staged else "k_parse_pairs",
                "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": measured_traffic(n, world) if staged else None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": parse_ms / args.steps,
                "index_kernel_ms": index_ms / args.steps,
                "index_algorithmic_bytes": BYTES_PER_REF_POS * r_bases,
                "note": "HBM is the contract's denominator; the largest kernel (k_hit_spec, 36 % of the step) is bound by the "
                        "integer pipe (LOP3/SHF at half rate: ALU pipe 90 % busy), the step issues 403 G warp-instructions "
                        "= 51 % of the chip's issue rate (DESIGN.md section 8, profiles/r02u_spec_probe_by_function.txt)"}

    # ---- parity, on EVERY rank: (1) a bounded sample of this rank's own rows against the reference CParser on the host
    # cores; (2) at 2 000 genomes every result of this rank against the reference's committed per-row digests
    parity = cpu = None
    if not args.no_parity:
        chk, kind = checker()
        cores = os.cpu_count() or 1
        threads = max(1, cores // world)
        n_rows = max(1, min(rows.size, args.cpu_rows or threads))
        per_row = max(1, min(n - 1, args.cpu_pairs // (n_rows * world)))
        crow = rows[np.linspace(0, rows.size - 1, n_rows).astype(np.int64)]
        off, q = sample_csr(n, crow, per_row, seed=77 + rank)
        barrier()
        ref_res = chk.run_rows(ss, None, rows=crow, row_off=off, q_ids=q, threads=threads)
        secs = chk.last_seconds
        got = m.run_rows(crow, off, q)
        ok = bool(np.array_equal(got, ref_res))
        full_rows = 0
        gold = golden_rowsums(n, ss)
        if gold is not None:
            sums, hsh = row_digest(out, rows.size)
            ok = ok and bool(np.array_equal(sums, gold[0][rows]) and np.array_equal(hsh, gold[1][rows]))
            full_rows = int(rows.size)
        checked = allsum(float(got.shape[0]))
        full_pairs = allsum(float(full_rows * (n - 1)))
        all_ok = allmin(1.0 if ok else 0.0) > 0.5
        busy_secs = allmax(secs)
        parity = {"ok": all_ok, "ranks": world, "pairs_checked_vs_reference_cpu": int(checked),
                  "pairs_checked_vs_reference_digests": int(full_pairs), "kind": kind,
                  "what": "every rank: a sample of its own rows re-run by the unmodified reference CParser"
                          + ("; every result of every rank against tests/golden/bench2000_rowsums.npz (reference row sums + position-weighted hashes of all 3 998 000 pairs)" if full_pairs else "")}
        if not all_ok:
            if not ok:
                print(f"PARITY FAILURE on rank {rank}", file=sys.stderr, flush=True)
            raise SystemExit("PARITY FAILURE at bench scale: CUDA results differ from the CPU reference")
        if not args.no_cpu_baseline:
            # all ranks ran their samples side by side on cores // world threads each: together one host-wide measurement
            cpu = {"value": checked / busy_secs, "unit": UNIT, "cores": threads * world, "kind": kind,
                   "sample": f"{world} x {n_rows} reference rows x {per_row} queries = {int(checked)} ordered pairs in {busy_secs:.1f} s "
                             f"({threads} threads per rank, ranks side by side); parity vs CUDA checked on the sample; built with {REF_BUILD}"}

    # ---- N >= 2: the in-process multi-GPU path (one library context, two devices, peer reads) on rank 0
    peer = None
    if world >= 2 and not args.no_parity:
        barrier()
        if rank == 0:
            names, seqs = synth.families(4, 6, 8000, 12000, seed=91, d_lo=0.02, d_hi=0.2)
            okp, n_sv = peer_group_check(hostio.SeqSet(names, seqs))
            peer = {"ok": okp, "surviving_pairs": n_sv}
            if not okp:
                raise SystemExit("PEER GROUP FAILURE: the two-device context differs from the one-device context")
        barrier()

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms_step, "higher_is_better": True, "scaling": args.scaling,
                "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": {"workload": f"synthetic viral all2all, {n} genomes ~50 kb in mutated families of 20 (seed {synth.SEED}), default params, {int(total_pairs)} ordered pairs per step, pair list sharded by reference row over {world} GPU(s)",
                           "genomes": n, "params": hostio.DEFAULT_PARAMS, "pairs_per_step": int(total_pairs),
                           "l2_policy": "inputs larger than L2: per step the kernels touch sequence records + indices of all references (~%.1f GB) and stream the hit records of every pair" % ((r_bases * 26 + total_len * 1.0) / 1e9)},
                "query_gbp_per_sec": total_qbases / (dev_ms_step / 1e3) / 1e9,
                "wall_ms_per_step": wall_ms_step, "parse_kernel_ms_per_step": parse_ms_step,
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
                "cpu_baseline": cpu, "parity": parity, "peer_filter_parity": (peer["ok"] if peer else None),
                "peer_group": peer, "result_checksum": checksum.tolist()}
        print(json.dumps(line), flush=True)
    m.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
