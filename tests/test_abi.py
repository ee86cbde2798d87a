"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol the header
declares, and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest

lz = importlib.import_module("lz-ani_b200")
ROOT = lz.ROOT


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "lzani_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(lzani_[a-z_0-9]+)\s*\(", hdr)))


def test_library_builds_and_exports_every_declared_symbol():
    lz.build()
    lib = C.CDLL(lz.LIB_PATH)
    syms = declared_symbols()
    assert {"lzani_create", "lzani_set_sequences", "lzani_run", "lzani_run_rows", "lzani_last_error",
            "lzani_destroy"} <= set(syms)
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/lzani_b200.h but not exported"
    assert lib.lzani_abi_version() == 1


def test_struct_layouts_match_header():
    # lzani_result_t == results_t: three int32 (defs.h:48-65); lzani_params_t: eight int32
    assert np.dtype(np.int32).itemsize * 3 == 12
    assert C.sizeof(lz.TimingStruct) == 6 * 8 + 5 * 8 + 2 * 4 + 3 * 8
    assert lz.PAIR_DTYPE.itemsize == 36          # lzani_pair_t: 2 x u32 + 2 x results_t + u32


def test_no_cpu_fallback_without_device():
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a CUDA device is present")
    except ImportError:
        pass
    with pytest.raises(lz.LzaniError, match="no CUDA device|CUDA"):
        lz.Matcher()


def test_product_sources_never_touch_the_oracle():
    # the product path (package + include) must not import, link or execute anything under oracle/
    bad = []
    for base in (os.path.join(ROOT, "lz-ani_b200"), os.path.join(ROOT, "include")):
        for dp, _, fns in os.walk(base):
            for fn in fns:
                if fn.endswith((".py", ".cu", ".h", ".cuh", ".cpp")):
                    txt = open(os.path.join(dp, fn), errors="ignore").read()
                    if re.search(r"oracle_lib|liblzani_oracle|libref_parser|lzani_oracle_|lzani_ref_|lzani_emul", txt):
                        bad.append(fn)
    assert not bad, bad


def test_shard_rows_partition():
    for n, w in [(10, 1), (10, 3), (7, 8), (2000, 8)]:
        parts = [lz.shard_rows(n, r, w) for r in range(w)]
        allr = np.sort(np.concatenate(parts))
        assert np.array_equal(allr, np.arange(n))
