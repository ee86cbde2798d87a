"""Host-side helpers that mirror the parts of the reference's loader the boundary depends on.

These are conveniences for tests, bench.py and Python users of the C ABI; they are NOT on the
hot path and are NOT a re-implementation of the reference's CLI/loader (which stays host C++,
see INTEGRATION.md).  Each function cites the reference code whose *format* it reproduces.
"""
from __future__ import annotations

import gzip
import os
from typing import Iterable, List, Sequence, Tuple

import numpy as np

# symbol codes, reference src/defs.h:24-30 and the dna_code table seq_reservoir.h:243-247:
# A,C,G,T (either case) -> 0..3, everything else -> 5 (code_N_seq)
_CODE = np.full(256, 5, dtype=np.uint8)
for _i, _c in enumerate("ACGT"):
    _CODE[ord(_c)] = _i
    _CODE[ord(_c.lower())] = _i

DEFAULT_PARAMS = dict(mal=11, msl=7, mrd=40, mqd=40, reg=35, aw=15, am=7, ar=3)  # params.h:38-45
PARAM_ORDER = ("mal", "msl", "mrd", "mqd", "reg", "aw", "am", "ar")


def params_array(params: dict | None = None) -> np.ndarray:
    p = dict(DEFAULT_PARAMS)
    if params:
        p.update(params)
    return np.array([p[k] for k in PARAM_ORDER], dtype=np.int32)


def encode(seq: bytes | str | np.ndarray) -> np.ndarray:
    """ASCII bases -> symbol codes 0..5 (seq_reservoir.h:243-247)."""
    if isinstance(seq, str):
        seq = seq.encode()
    if isinstance(seq, (bytes, bytearray)):
        seq = np.frombuffer(bytes(seq), dtype=np.uint8)
    return _CODE[seq]


def pack3(codes: np.ndarray) -> np.ndarray:
    """Symbol codes -> the reservoir's `three_in_byte` packing (seq_reservoir.cpp:39-57):
    byte = 36a + 6b + c; a trailing group of 1 or 2 symbols is high-aligned."""
    codes = np.asarray(codes, dtype=np.uint8)
    n = codes.size
    full = n // 3
    out = np.zeros((n + 2) // 3, dtype=np.uint8)
    if full:
        t = codes[: 3 * full].reshape(full, 3).astype(np.uint16)
        out[:full] = (36 * t[:, 0] + 6 * t[:, 1] + t[:, 2]).astype(np.uint8)
    rem = n - 3 * full
    if rem == 2:
        out[full] = 36 * int(codes[-2]) + 6 * int(codes[-1])
    elif rem == 1:
        out[full] = 36 * int(codes[-1])
    return out


def unpack3(packed: np.ndarray, n: int) -> np.ndarray:
    """Inverse of pack3 (seq_view::unpack3, seq_reservoir.h:145-169)."""
    packed = np.asarray(packed, dtype=np.uint8).astype(np.uint16)
    out = np.empty(3 * packed.size, dtype=np.uint8)
    out[0::3] = packed // 36
    out[1::3] = (packed // 6) % 6
    out[2::3] = packed % 6
    return out[:n]


def read_fasta(path: str) -> List[Tuple[str, bytes]]:
    """Multi-FASTA records; name = header up to the first space (seq_reservoir.cpp:77-81,
    183-196).  Plain or gzip."""
    opener = gzip.open if path.endswith(".gz") else open
    recs: List[Tuple[str, bytes]] = []
    name = None
    chunks: List[bytes] = []
    with opener(path, "rb") as f:
        for line in f:
            line = line.rstrip(b"\r\n")
            if line.startswith(b">"):
                if name is not None:
                    recs.append((name, b"".join(chunks)))
                name = line[1:].split(b" ")[0].decode()
                chunks = []
            elif name is not None:
                chunks.append(line)
    if name is not None:
        recs.append((name, b"".join(chunks)))
    return recs


def read_fasta_dir(dirname: str) -> List[Tuple[str, bytes]]:
    """All regular files of a directory as multi-FASTA, in sorted path order.  (The reference
    iterates `directory_iterator` in OS order; the later length-descending reorder makes the
    final ids independent of it except for equal (length, name) keys.)"""
    recs: List[Tuple[str, bytes]] = []
    for fn in sorted(os.listdir(dirname)):
        p = os.path.join(dirname, fn)
        if os.path.isfile(p):
            recs.extend(read_fasta(p))
    return recs


def reorder(names: Sequence[str], lens: Sequence[int]) -> np.ndarray:
    """Permutation of CSeqReservoir::reorder_items (seq_reservoir.cpp:215-251) for
    no_parts == 1: stable sort by (len - 2) descending, then name ascending.
    Returns `order` such that new id i is old id order[i]."""
    idx = list(range(len(names)))
    idx.sort(key=lambda i: (-(int(lens[i]) - 2), names[i]))
    return np.array(idx, dtype=np.int64)


class SeqSet:
    """A set of sequences in boundary form: three_in_byte bytes + lengths (+ names)."""

    def __init__(self, names: Sequence[str], codes: Iterable[np.ndarray]):
        self.names = list(names)
        self.codes = [np.ascontiguousarray(c, dtype=np.uint8) for c in codes]
        self.lens = np.array([c.size for c in self.codes], dtype=np.uint32)
        self.packed = [pack3(c) for c in self.codes]

    @classmethod
    def from_records(cls, recs: Sequence[Tuple[str, bytes]], reorder_items: bool = True) -> "SeqSet":
        names = [r[0] for r in recs]
        codes = [encode(r[1]) for r in recs]
        if reorder_items:
            order = reorder(names, [c.size for c in codes])
            names = [names[i] for i in order]
            codes = [codes[i] for i in order]
        return cls(names, codes)

    def __len__(self) -> int:
        return len(self.names)

    def subset(self, ids: Sequence[int]) -> "SeqSet":
        return SeqSet([self.names[i] for i in ids], [self.codes[i] for i in ids])

    def total_bases(self) -> int:
        return int(self.lens.sum())


def all2all_csr(n: int, rows: Sequence[int] | None = None):
    """CSR of the dense all2all pair list (every q != r, increasing q), the order
    do_matching visits them in (lz_matcher.cpp:216-232)."""
    rows = list(range(n)) if rows is None else list(rows)
    off = np.zeros(len(rows) + 1, dtype=np.uint64)
    q = np.empty(len(rows) * max(n - 1, 0), dtype=np.uint32)
    k = 0
    for j, r in enumerate(rows):
        ids = np.concatenate([np.arange(0, r, dtype=np.uint32), np.arange(r + 1, n, dtype=np.uint32)])
        q[k : k + ids.size] = ids
        k += ids.size
        off[j + 1] = k
    return np.array(rows, dtype=np.uint32), off, q
