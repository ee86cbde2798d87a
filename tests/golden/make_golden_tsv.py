#!/usr/bin/env python
"""Golden TSV files for the CLI-level drop-in test (tests/test_integration_cli.py), produced by the
UNMODIFIED reference CLI (oracle/_ref/lz-ani, built from /root/reference by oracle/Makefile).
Run in the build container only:  python tests/golden/make_golden_tsv.py

For the two cases the reference ships goldens for, the output is first checked to be byte-identical
to the reference's own fixture (test/vir61.ani.tsv, example/output/ani.tsv)."""
import gzip
import hashlib
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
CLI = os.path.join(ROOT, "oracle", "_ref", "lz-ani")

ALT = "--mal 15 --msl 9 --mrd 60 --mqd 60 --reg 60 --aw 20 --am 9 --ar 4".split()
# name: (input args relative to REF, extra args, reference fixture to compare with or None)
CASES = {
    "vir61_default": (["--in-dir", "test/vir61"], [], "test/vir61.ani.tsv"),
    "vir61_alt_complete": (["--in-dir", "test/vir61"], ALT + ["--out-format", "complete"], None),
    "example_default": (["--in-fasta", "example/multifasta.fna"], [], "example/output/ani.tsv"),
    "example_filter09": (["--in-fasta", "example/multifasta.fna"], ["--flt-kmerdb", "example/fltr.txt", "0.9"], None),
    # --out-filter (lz_matcher.cpp:451-462): the device-side pre-filter of lzani_filter_pairs must keep exactly these lines
    "vir61_outfilter": (["--in-dir", "test/vir61"], ["--out-filter", "ani", "0.9", "--out-filter", "qcov", "0.3"], None),
    "example_outfilter": (["--in-fasta", "example/multifasta.fna"],
                          ["--out-filter", "tani", "0.85", "--out-filter", "rcov", "0.9", "--out-filter", "gani", "0.8"], None),
    "example_filter09_outfilter": (["--in-fasta", "example/multifasta.fna"],
                                   ["--flt-kmerdb", "example/fltr.txt", "0.9", "--out-filter", "ani", "0.97"], None),
    # submits without pairs: a kmer-db filter nothing passes, a set of one sequence -- header-only output
    "example_filter_none": (["--in-fasta", "example/multifasta.fna"], ["--flt-kmerdb", "example/fltr.txt", "1.5"], None),
    "single_default": (["--in-fasta", "@first:example/multifasta.fna"], [], None),
}


def first_record(path, tmp):
    """A FASTA holding only the first record of `path`."""
    out = os.path.join(tmp, "first.fna")
    with open(path) as f, open(out, "w") as o:
        n = 0
        for line in f:
            if line.startswith(">"):
                n += 1
                if n > 1:
                    break
            o.write(line)
    return out


def main():
    tmp = tempfile.mkdtemp()
    only = set(sys.argv[1:])
    for name, (inp, extra, fixture) in CASES.items():
        if only and name not in only:
            continue
        inp = [first_record(os.path.join(REF, a[7:]), tmp) if a.startswith("@first:") else a for a in inp]
        out = os.path.join(tmp, name + ".tsv")
        args = [a if not (a.startswith("test/") or a.startswith("example/")) else os.path.join(REF, a) for a in inp + extra]
        subprocess.run([CLI, "all2all"] + args + ["--out", out, "-V", "1"], check=True, stderr=subprocess.DEVNULL)
        data = open(out, "rb").read()
        if fixture:
            want = open(os.path.join(REF, fixture), "rb").read()
            assert data == want, f"{name}: reference CLI output differs from the reference's own golden {fixture}"
        with gzip.GzipFile(os.path.join(HERE, name + ".ani.tsv.gz"), "wb", mtime=0) as f:
            f.write(data)
        ids = open(out.replace(".tsv", ".ids.tsv"), "rb").read() if os.path.exists(out.replace(".tsv", ".ids.tsv")) else b""
        with gzip.GzipFile(os.path.join(HERE, name + ".ani.ids.tsv.gz"), "wb", mtime=0) as f:
            f.write(ids)
        print(name, len(data), hashlib.md5(data).hexdigest())
    if only:
        shutil.rmtree(tmp)
        return
    # --out-alignment (line order depends on thread interleaving in the reference: compare sorted)
    out = os.path.join(tmp, "aln_main.tsv")
    aln = os.path.join(tmp, "aln.tsv")
    subprocess.run([CLI, "all2all", "--in-fasta", os.path.join(REF, "example/multifasta.fna"), "--out", out,
                    "--out-alignment", aln, "-V", "1"], check=True, stderr=subprocess.DEVNULL)
    lines = sorted(open(aln, "rb").read().splitlines(keepends=True))
    want = sorted(open(os.path.join(REF, "example/output/ani.aln.tsv"), "rb").read().splitlines(keepends=True))
    assert lines == want, "reference CLI alignment output differs from the reference's own golden"
    with gzip.GzipFile(os.path.join(HERE, "example_default.aln.sorted.tsv.gz"), "wb", mtime=0) as f:
        f.write(b"".join(lines))
    print("example alignment", len(lines), "lines")
    shutil.copy(os.path.join(REF, "example", "fltr.txt"), os.path.join(HERE, "example_fltr.txt"))
    shutil.rmtree(tmp)


if __name__ == "__main__":
    sys.exit(main())
