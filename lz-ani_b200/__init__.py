"""lz-ani_b200 -- B200-native LZ-ANI pair-comparison hot path (host-side Python binding).

The product is `liblzani_b200.so` (hand-written sm_100a CUDA behind the C ABI of
include/lzani_b200.h).  This module is the thin ctypes mirror of that ABI plus the small
host helpers tests and bench.py need (hostio: three_in_byte packing, FASTA, reorder; synth:
seeded synthetic genome sets).  There is no CPU fallback: if the shared library is missing
or no CUDA device is present, constructing `Matcher` raises.

Import with importlib because of the hyphen in the directory name:

    lz = importlib.import_module("lz-ani_b200")
    m = lz.Matcher(params=None, device=0)
    m.set_sequences(seqset)              # hostio.SeqSet
    res = m.all2all()                    # (n*(n-1), 3) int32: nt_match, nt_mismatch, num_alns
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import hostio, sharding, synth  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
# LZANI_LIB: an alternative build of the same library (kernel tuning experiments, tools/variants.sh)
LIB_PATH = os.environ.get("LZANI_LIB") or os.path.join(_HERE, "liblzani_b200.so")
CSRC = os.path.join(_HERE, "csrc")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA extension in-tree for sm_100a (cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in ("kernels.cu", "postproc.cu", "pairlist.cu", "capi.cu", "noalloc.cu")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("kernels.h", "layout.h", "parse_core.h", "lane_walk.h", "regions.h", "simt.h", "outfilter.h")] + [
        os.path.join(ROOT, "include", "lzani_b200.h")]
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + srcs + ["-o", LIB_PATH]
    subprocess.run(cmd, check=True)
    return LIB_PATH


class TimingStruct(C.Structure):
    _fields_ = [("h2d_ms", C.c_double), ("pack_ms", C.c_double), ("index_ms", C.c_double), ("parse_ms", C.c_double),
                ("d2h_ms", C.c_double), ("total_ms", C.c_double), ("n_pairs", C.c_uint64), ("q_bases", C.c_uint64),
                ("r_bases", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("launches", C.c_uint32), ("n_batches", C.c_uint32),
                ("filter_ms", C.c_double), ("filter_d2h_bytes", C.c_uint64), ("pairlist_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class LzaniError(RuntimeError):
    pass


_lib = None


def load_library() -> C.CDLL:
    """dlopen the in-tree CUDA library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LzaniError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                         "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    lib.lzani_abi_version.restype = C.c_int
    lib.lzani_create.argtypes = [C.POINTER(vp), vp, vp, C.c_int]
    lib.lzani_set_sequences.argtypes = [vp, u32, vp, vp]
    lib.lzani_set_sequences_text.argtypes = [vp, u32, vp, vp]
    lib.lzani_run.argtypes = [vp, u32, vp, vp, vp]
    lib.lzani_run_rows.argtypes = [vp, u32, vp, vp, vp, vp]
    lib.lzani_run_rows_resident.argtypes = [vp, u32, vp, vp, vp, C.c_int, vp]
    lib.lzani_run_rows_regions.argtypes = [vp, u32, vp, vp, vp, vp, vp, vp, u64, vp]
    lib.lzani_filter_pairs.argtypes = [vp, vp, C.c_int, vp, vp, vp, u64, vp]
    lib.lzani_build_pair_list.argtypes = [vp, u32, u64, vp, vp, vp, vp, vp, u64, vp]
    lib.lzani_filter_dense_noalloc.argtypes = [C.c_int, u32, vp, vp, vp, vp, u64, vp, vp, C.c_char_p, C.c_size_t]
    lib.lzani_get_timing.argtypes = [vp, C.POINTER(TimingStruct)]
    lib.lzani_set_index_budget.argtypes = [vp, u64]
    lib.lzani_last_error.argtypes = [vp]
    lib.lzani_last_error.restype = C.c_char_p
    lib.lzani_destroy.argtypes = [vp]
    lib.lzani_destroy.restype = None
    _lib = lib
    return lib


class Matcher:
    """Mirror of what CLZMatcher::do_matching does with its CParser objects
    (reference src/lz_matcher.cpp:172-277), batch form."""

    def __init__(self, params: dict | None = None, device: int | None = None, devices=None):
        """device: one GPU.  devices=[d0, d1, ...]: ONE context driving several GPUs -- the library owns the devices
        and deals the reference rows over them (lzani_create with n_devices > 1)."""
        self.lib = load_library()
        self.params = hostio.params_array(params)
        self.h = C.c_void_p()
        dev, n_dev = None, 0
        if devices is not None:
            dev, n_dev = (C.c_int * len(devices))(*[int(d) for d in devices]), len(devices)
        elif device is not None:
            dev, n_dev = (C.c_int * 1)(int(device)), 1
        rc = self.lib.lzani_create(C.byref(self.h), self.params.ctypes.data, dev, n_dev)
        if rc != 0:
            msg = self.lib.lzani_last_error(self.h).decode() if self.h else "allocation failed"
            if self.h:
                self.lib.lzani_destroy(self.h)
                self.h = None
            raise LzaniError(f"lzani_create failed ({rc}): {msg}")
        self.n = 0

    def _check(self, rc: int, what: str):
        if rc != 0:
            raise LzaniError(f"{what} failed ({rc}): {self.lib.lzani_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.lzani_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_index_budget(self, nbytes: int):
        self._check(self.lib.lzani_set_index_budget(self.h, int(nbytes)), "lzani_set_index_budget")

    def set_sequences(self, ss: "hostio.SeqSet"):
        n = len(ss)
        ptrs = (C.c_void_p * max(n, 1))()
        for i, p in enumerate(ss.packed):
            ptrs[i] = p.ctypes.data
        lens = np.ascontiguousarray(ss.lens, dtype=np.uint32)
        self._check(self.lib.lzani_set_sequences(self.h, n, ptrs, lens.ctypes.data), "lzani_set_sequences")
        self.n = n

    def set_sequences_text(self, texts):
        """Device-side ingest (lzani_set_sequences_text): texts[i] = bytes / uint8 array, one byte per base,
        as CSeqReservoir::append receives the sequence (seq_reservoir.cpp:18-87)."""
        arrs = [np.frombuffer(t, dtype=np.uint8) if isinstance(t, (bytes, bytearray)) else np.ascontiguousarray(t, dtype=np.uint8)
                for t in texts]
        n = len(arrs)
        ptrs = (C.c_void_p * max(n, 1))()
        for i, a in enumerate(arrs):
            ptrs[i] = a.ctypes.data if a.size else None
        lens = np.array([a.size for a in arrs], dtype=np.uint32)
        self._check(self.lib.lzani_set_sequences_text(self.h, n, ptrs, lens.ctypes.data), "lzani_set_sequences_text")
        self.n = n

    def timing(self) -> dict:
        t = TimingStruct()
        self._check(self.lib.lzani_get_timing(self.h, C.byref(t)), "lzani_get_timing")
        return t.as_dict()

    def all2all(self) -> np.ndarray:
        out = np.zeros((self.n * max(self.n - 1, 0), 3), dtype=np.int32)
        self._check(self.lib.lzani_run(self.h, self.n, None, None, out.ctypes.data), "lzani_run")
        return out

    def run(self, row_off, q_ids) -> np.ndarray:
        row_off = np.ascontiguousarray(row_off, dtype=np.uint64)
        q_ids = np.ascontiguousarray(q_ids, dtype=np.uint32)
        out = np.zeros((int(row_off[-1] - row_off[0]), 3), dtype=np.int32)
        self._check(self.lib.lzani_run(self.h, self.n, row_off.ctypes.data, q_ids.ctypes.data, out.ctypes.data), "lzani_run")
        return out

    def run_rows(self, rows, row_off=None, q_ids=None, out: np.ndarray | None = None, resident_repeat: int = 0):
        """rows: reference ids; (row_off, q_ids) CSR over those rows or None for every q != r.
        resident_repeat > 0 -> lzani_run_rows_resident (kernels only, `repeat` passes)."""
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        if q_ids is None:
            n_pairs = rows.size * max(self.n - 1, 0)
            off_p = q_p = None
        else:
            row_off = np.ascontiguousarray(row_off, dtype=np.uint64)
            q_ids = np.ascontiguousarray(q_ids, dtype=np.uint32)
            n_pairs = int(row_off[-1] - row_off[0])
            off_p, q_p = row_off.ctypes.data, q_ids.ctypes.data
        if out is None:
            out = np.zeros((n_pairs, 3), dtype=np.int32)
        assert out.shape == (n_pairs, 3) and out.dtype == np.int32 and out.flags.c_contiguous
        if resident_repeat > 0:
            self._check(self.lib.lzani_run_rows_resident(self.h, rows.size, rows.ctypes.data, off_p, q_p,
                                                         int(resident_repeat), out.ctypes.data), "lzani_run_rows_resident")
        else:
            self._check(self.lib.lzani_run_rows(self.h, rows.size, rows.ctypes.data, off_p, q_p, out.ctypes.data),
                        "lzani_run_rows")
        return out


def _run_rows_regions(self, rows, row_off=None, q_ids=None, cap: int = 1 << 16):
    """Counts plus the kept regions of every pair (what --out-alignment needs; replaces
    CParser::get_parsing, parser.cpp:786-837).  Returns (out, region_off, regions[n, 6]) with the
    region columns ref_start, ref_end, seq_start, seq_end, num_matches, num_mismatches."""
    rows = np.ascontiguousarray(rows, dtype=np.uint32)
    if q_ids is None:
        n_pairs = rows.size * max(self.n - 1, 0)
        off_p = q_p = None
    else:
        row_off = np.ascontiguousarray(row_off, dtype=np.uint64)
        q_ids = np.ascontiguousarray(q_ids, dtype=np.uint32)
        n_pairs = int(row_off[-1] - row_off[0])
        off_p, q_p = row_off.ctypes.data, q_ids.ctypes.data
    out = np.zeros((n_pairs, 3), dtype=np.int32)
    region_off = np.zeros(n_pairs + 1, dtype=np.uint64)
    for _ in range(2):
        regions = np.zeros((max(cap, 1), 6), dtype=np.int32)
        n_reg = C.c_uint64(0)
        rc = self.lib.lzani_run_rows_regions(self.h, rows.size, rows.ctypes.data, off_p, q_p, out.ctypes.data,
                                             region_off.ctypes.data, regions.ctypes.data, cap, C.byref(n_reg))
        if rc == 4 and n_reg.value > cap:      # LZANI_ERR_NOMEM: the library says how many it needs
            cap = int(n_reg.value)
            continue
        self._check(rc, "lzani_run_rows_regions")
        return out, region_off, regions[: int(n_reg.value)].copy()
    raise LzaniError("lzani_run_rows_regions: region count changed between calls")


Matcher.run_rows_regions = _run_rows_regions

PAIR_DTYPE = np.dtype([("a", np.uint32), ("b", np.uint32), ("ab", np.int32, 3), ("ba", np.int32, 3), ("lines", np.uint32)])
FILTER_KEYS = ("gani", "ani", "tani", "qcov", "rcov")


def _filter_pairs(self, eff_len, flt: dict, group=None, cap: int = 1 << 12):
    """Device-side pairing + derived metrics + --out-filter compaction (lzani_filter_pairs;
    replaces the arithmetic of store_results, lz_matcher.cpp:404-463) over the results the last
    run_rows(..., resident_repeat=1) left on the device.  flt: thresholds by name (missing = 0).
    group: the Matchers (one per GPU) that together hold both directions; None = this one.
    Returns a structured array (PAIR_DTYPE) of the surviving unordered pairs in CSR order."""
    eff_len = np.ascontiguousarray(eff_len, dtype=np.uint32)
    assert eff_len.size == self.n
    thr = np.array([float(flt.get(k, 0.0)) for k in FILTER_KEYS], dtype=np.float64)
    grp_p, n_grp = None, 0
    if group:
        grp_p = (C.c_void_p * len(group))(*[g.h for g in group])
        n_grp = len(group)
    for _ in range(2):
        pairs = np.zeros(max(cap, 1), dtype=PAIR_DTYPE)
        n_out = C.c_uint64(0)
        rc = self.lib.lzani_filter_pairs(self.h, grp_p, n_grp, eff_len.ctypes.data, thr.ctypes.data,
                                         pairs.ctypes.data, cap, C.byref(n_out))
        if rc == 4 and n_out.value > cap:
            cap = int(n_out.value)
            continue
        self._check(rc, "lzani_filter_pairs")
        return pairs[: int(n_out.value)].copy()
    raise LzaniError("lzani_filter_pairs: survivor count changed between calls")


Matcher.filter_pairs = _filter_pairs


def _build_pair_list(self, n: int, edge_i, edge_j, reorder_map=None):
    """Sparse pair list on the device (lzani_build_pair_list; replaces the symmetric closure and id
    remap of CFilter, filter.cpp:243-345).  Returns (row_off[n+1] uint64, q_ids uint32)."""
    edge_i = np.ascontiguousarray(edge_i, dtype=np.uint32)
    edge_j = np.ascontiguousarray(edge_j, dtype=np.uint32)
    assert edge_i.size == edge_j.size
    mp = None
    if reorder_map is not None:
        reorder_map = np.ascontiguousarray(reorder_map, dtype=np.uint32)
        assert reorder_map.size == n
        mp = reorder_map.ctypes.data
    row_off = np.zeros(n + 1, dtype=np.uint64)
    q_ids = np.zeros(max(2 * edge_i.size, 1), dtype=np.uint32)
    n_ent = C.c_uint64(0)
    self._check(self.lib.lzani_build_pair_list(self.h, n, edge_i.size, edge_i.ctypes.data, edge_j.ctypes.data, mp,
                                               row_off.ctypes.data, q_ids.ctypes.data, q_ids.size, C.byref(n_ent)),
                "lzani_build_pair_list")
    return row_off, q_ids[: int(n_ent.value)].copy()


Matcher.build_pair_list = _build_pair_list

from .sharding import shard_rows  # noqa: E402,F401
