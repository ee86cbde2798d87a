"""ctypes front-ends of the two CPU checkers (TEST INFRASTRUCTURE ONLY):

* `Oracle`  -- oracle/liblzani_oracle.so, our plain-C restatement (always buildable);
* `RefShim` -- oracle/_ref/libref_parser.so, the UNMODIFIED reference CParser behind
               oracle/ref_shim.cpp (exists where /root/reference was available at build time).

Both expose run_rows(seqset, params, rows=None, row_off=None, q_ids=None, threads=N) -> (n_pairs, 3)
int32 array in CSR order, and regions(seqset, params, r, q).
"""
from __future__ import annotations

import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
hostio = importlib.import_module("lz-ani_b200.hostio")

ORACLE_SO = os.path.join(ROOT, "oracle", "liblzani_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_parser.so")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "lz-ani")


def build_oracle(want_ref: bool = True) -> None:
    targets = ["oracle"] + (["ref"] if want_ref else [])
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")] + targets, check=True)


def _ptr_array(packed):
    arr = (C.c_void_p * len(packed))()
    for i, p in enumerate(packed):
        arr[i] = p.ctypes.data
    return arr


class _Checker:
    def __init__(self, so_path: str, prefix: str):
        self.lib = C.CDLL(so_path)
        self._run = getattr(self.lib, prefix + "_run_rows")
        self._run.restype = C.c_double
        self._run.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        self._regions = getattr(self.lib, prefix + "_regions")
        self._regions.restype = C.c_int
        self._regions.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p,
                                  C.c_uint32, C.c_void_p]
        self.last_seconds = 0.0

    def run_rows(self, ss, params=None, rows=None, row_off=None, q_ids=None, threads=0):
        p8 = hostio.params_array(params)
        n = len(ss)
        ptrs = _ptr_array(ss.packed)
        lens = np.ascontiguousarray(ss.lens, dtype=np.uint32)
        if rows is None:
            n_rows, rows_p = n, None
        else:
            rows = np.ascontiguousarray(rows, dtype=np.uint32)
            n_rows, rows_p = rows.size, rows.ctypes.data
        if q_ids is None:
            n_pairs = n_rows * max(n - 1, 0)
            off_p = q_p = None
        else:
            row_off = np.ascontiguousarray(row_off, dtype=np.uint64)
            q_ids = np.ascontiguousarray(q_ids, dtype=np.uint32)
            n_pairs = int(row_off[-1])
            off_p, q_p = row_off.ctypes.data, q_ids.ctypes.data
        out = np.zeros((n_pairs, 3), dtype=np.int32)
        if threads <= 0:
            threads = os.cpu_count() or 1
        secs = self._run(p8.ctypes.data, n, ptrs, lens.ctypes.data, n_rows, rows_p, off_p, q_p,
                         out.ctypes.data, int(threads))
        if secs < 0:
            raise RuntimeError("checker rejected its arguments")
        self.last_seconds = secs
        return out

    def regions(self, ss, params, r, q, cap=1 << 16):
        p8 = hostio.params_array(params)
        out = np.zeros((cap, 6), dtype=np.int32)
        st = np.zeros(3, dtype=np.int32)
        n = self._regions(p8.ctypes.data, ss.packed[r].ctypes.data, int(ss.lens[r]), ss.packed[q].ctypes.data,
                          int(ss.lens[q]), out.ctypes.data, cap, st.ctypes.data)
        return out[:n].copy(), st


class Oracle(_Checker):
    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build_oracle(want_ref=False)
        super().__init__(ORACLE_SO, "lzani_oracle")


class RefShim(_Checker):
    def __init__(self):
        super().__init__(REF_SO, "lzani_ref")

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)


EMUL_SO = os.path.join(ROOT, "tests", "emul", "liblzani_emul.so")


class Emul(_Checker):
    """Host emulation of the CUDA kernel SOURCE (tests/emul/emul.cpp) -- test only."""

    def __init__(self):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests", "emul")], check=True)
        super().__init__(EMUL_SO, "lzani_emul")

    def set_staged(self, on):
        """0 / False: the one-kernel form (the warp builds its batches itself); 1 / True: the staged pipeline
        (hit records computed ahead per query position) with the warp-per-pair walk; 2: the staged pipeline with
        the lane-per-pair walk (lane_walk.h, the product's default)."""
        self.lib.lzani_emul_set_staged(int(on))
