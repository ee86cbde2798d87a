#!/usr/bin/env python
"""Differential fuzz of the CUDA library against the oracle (the generators of tests/test_gpu_parity.py::
test_cuda_random_parameters_and_ragged_sets_differential over many more seeds).

    python tools/gpu_fuzz.py <first_seed> <n_seeds>
"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")]
lz = importlib.import_module("lz-ani_b200")
from oracle_lib import Oracle  # noqa: E402  (checker only)
from test_emul import _weird_sequence  # noqa: E402

hostio, synth = lz.hostio, lz.synth
first, count = int(sys.argv[1]), int(sys.argv[2])
orc = Oracle()
t0, bad, pairs, kinds = time.time(), 0, 0, [0, 0, 0]
for seed in range(first, first + count):
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
    kinds[seed % 3] += 1
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
    m = lz.Matcher(params=params, device=0)
    try:
        m.set_sequences(ss)
        got = m.all2all()
    finally:
        m.close()
    want = orc.run_rows(ss, params, threads=8)
    pairs += len(want)
    if not np.array_equal(got, want):
        bad += 1
        print("MISMATCH seed", seed, params, flush=True)
print(f"seeds {first}..{first + count - 1}: {count} sets ({kinds[0]} default parameters, {kinds[1]} random windows, {kinds[2]} random "
      f"parameter sets), {pairs} ordered pairs, {bad} sets with a difference, {time.time() - t0:.1f} s")
sys.exit(1 if bad else 0)
