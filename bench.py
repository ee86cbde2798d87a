#!/usr/bin/env python
"""bench.py -- the LZ-ANI all2all pair-comparison hot path on B200 (contract: see DESIGN.md).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--genomes G]

Workload (BASELINE.json configs[1]): seeded synthetic viral genomes, ~50 kb, families of 20
with 1-30 % divergence + indels + block events (lz-ani_b200/synth.py, seed 20261104), all2all,
default parameters.  One STEP = one pass of the hot path over the whole pair list of this rank:
reference index build (K2) + the staged parse (K3: per chunk of pairs the data-parallel scan / expand /
probe / spec kernels and the warp-per-pair walk; LZANI_STAGED=0 selects the one-kernel form).  N=1: G = 2000 genomes -> 3 998 000
ordered pairs per step.  N>1 (one process per GPU, torchrun): WEAK scaling -- the set grows to
G_N = round(G*sqrt(N)) genomes so that every rank keeps ~G*(G-1) pairs; rank r owns reference
rows r, r+N, ... (pair list sharded by reference block, sequences replicated, no collective
on the data path).

  value : ordered pairs/s, whole job, sequences + pair list resident in HBM, device-timed
          (CUDA events on the library's stream), max over ranks.
  e2e   : the same metric through the public C ABI with HOST buffers each step:
          lzani_set_sequences (H2D + K1 pack) + lzani_run_rows (pair list H2D, K2, K3,
          results D2H), wall-clock around the calls, max over ranks.
  cpu_baseline / --impl reference : the UNMODIFIED reference CParser (oracle/_ref, built from
          /root/reference by oracle/Makefile) on the box's host cores, bounded sample.
"""
from __future__ import annotations

import argparse
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


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_traffic(n_genomes: int, world: int):
    """DRAM bytes of the parse stage per step from the committed ncu capture of THIS workload
    (profiles/traffic.json, written from `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`);
    None when no capture matches the configuration."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        t = json.load(open(p))
        if int(t["genomes"]) == n_genomes and world == 1:
            return float(t["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


def make_set(n_genomes: int):
    cache = f"/tmp/lzani_synth_{synth.SEED}_{n_genomes}.npz"
    if os.path.exists(cache):
        z = np.load(cache)
        lens = z["lens"]
        flat = z["flat"]
        offs = np.concatenate([[0], np.cumsum(lens)])
        seqs = [flat[offs[i]:offs[i + 1]] for i in range(len(lens))]
        names = [str(x) for x in z["names"]]
    else:
        names, seqs = synth.viral(n_genomes)
        try:
            np.savez(cache, names=np.array(names), lens=np.array([s.size for s in seqs]), flat=np.concatenate(seqs))
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


def reference_arm(args, rank, world):
    """The reference's own CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle, RefShim
    cores = os.cpu_count() or 1
    G = int(round(args.genomes * math.sqrt(args.gpus)))
    ss = make_set(G)
    n = len(ss)
    if RefShim.available():
        chk, kind = RefShim(), "reference"
    else:
        chk, kind = Oracle(), "port"
    n_rows = max(1, min(n, args.cpu_rows or cores))
    rows_all = np.linspace(0, n - 1, n_rows * (args.warmup + args.steps)).astype(np.uint32)
    times, pairs = [], n_rows * (n - 1)
    for s in range(args.warmup + args.steps):
        rows = rows_all[s * n_rows:(s + 1) * n_rows]
        chk.run_rows(ss, None, rows=rows, threads=cores)
        if s >= args.warmup:
            times.append(chk.last_seconds)
    ms = 1e3 * float(np.mean(times))
    v = pairs / (ms / 1e3)
    qb = float(ss.lens.sum()) / n * pairs
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"synthetic viral all2all, {n} genomes ~50 kb, default params; each step = {n_rows} reference rows x {n - 1} queries ({pairs} ordered pairs) on {cores} host threads",
                       "genomes": n, "params": hostio.DEFAULT_PARAMS},
            "query_gbp_per_sec": qb / (ms / 1e3) / 1e9,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{n_rows} reference rows x {n - 1} queries per step"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--genomes", type=int, default=2000, help="genomes at N=1 (BASELINE configs[1]: 2000)")
    ap.add_argument("--cpu-rows", type=int, default=0, help="reference rows per CPU-baseline step (default: host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
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

    def allmax(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    G = int(round(args.genomes * math.sqrt(world)))
    ss = make_set(G)
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
    checksum = out.astype(np.int64).sum(axis=0).tolist()

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
        assert out.astype(np.int64).sum(axis=0).tolist() == checksum, "e2e results differ from the resident run"

    # ---- roofline of the parse stage, per step.  The parse is a pipeline of kernels per chunk of pairs
    # (scan, expand, probe, spec, walk; k_walk_pairs is the dominant one, see profiles/*_launch_summary.txt);
    # SURVEY 8d's algorithmic bytes (4.375 B per query base + 12 B per pair) are those of the WHOLE parse of a
    # pair, so they are divided by the time of the whole stage (CUDA events around it on the library's
    # stream), not by the walk kernel's share of it.
    peak, peak_src = load_peaks()
    alg_bytes = BYTES_PER_QUERY_BASE * my_qbases + 12.0 * my_pairs
    achieved = alg_bytes / (parse_ms / args.steps / 1e3) / 1e9 if parse_ms > 0 else 0.0
    staged = os.environ.get("LZANI_STAGED", "1") != "0"
    roofline = {"bound": "hbm", "kernel": "parse stage: k_scan_matrix + k_expand_hits + k_hit_probe + k_hit_spec + k_walk_pairs (dominant)" if staged else "k_parse_pairs",
                "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": measured_traffic(n, world) if staged else None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": parse_ms / args.steps,
                "index_kernel_ms": index_ms / args.steps,
                "index_algorithmic_bytes": BYTES_PER_REF_POS * r_bases}

    # ---- CPU baseline beside it (rank 0, N=1 only): the reference CParser on the host cores
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from oracle_lib import Oracle, RefShim   # checker / baseline only, never on the product path
        cores = os.cpu_count() or 1
        chk, kind = (RefShim(), "reference") if RefShim.available() else (Oracle(), "port")
        n_rows = max(1, min(n, args.cpu_rows or cores))
        crow = np.linspace(0, n - 1, n_rows).astype(np.uint32)
        ref_res = chk.run_rows(ss, None, rows=crow, threads=cores)
        secs = chk.last_seconds
        # the sample doubles as a parity check at bench scale
        got = m.run_rows(crow)
        if not np.array_equal(got, ref_res):
            raise SystemExit("PARITY FAILURE at bench scale: CUDA results differ from the CPU reference")
        cpu = {"value": n_rows * (n - 1) / secs, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"{n_rows} evenly spaced reference rows x {n - 1} queries = {n_rows * (n - 1)} ordered pairs in {secs:.1f} s; parity vs CUDA checked on the sample"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": {"workload": f"synthetic viral all2all, {n} genomes ~50 kb in mutated families of 20 (seed {synth.SEED}), default params, {int(total_pairs)} ordered pairs per step, pair list sharded by reference row over {world} GPU(s)",
                           "genomes": n, "params": hostio.DEFAULT_PARAMS, "pairs_per_step": int(total_pairs),
                           "l2_policy": "inputs larger than L2: per step the kernels touch sequence records + indices of all references (~%.1f GB)" % ((r_bases * 26 + total_len * 1.0) / 1e9)},
                "query_gbp_per_sec": total_qbases / (dev_ms_step / 1e3) / 1e9,
                "wall_ms_per_step": wall_ms_step, "parse_kernel_ms_per_step": parse_ms_step,
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
                "cpu_baseline": cpu, "result_checksum": checksum}
        print(json.dumps(line), flush=True)
    m.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
