"""CLI-level drop-in test: the reference's own host program (CLI, loader, filter, reorder, TSV
writers -- compiled unmodified by integration/Makefile) with CLZMatcher::do_matching replaced by
the B200 library must write TSV files byte-identical to the reference's
(reference CI: `cmp ani.tsv test/vir61.ani.tsv`, .github/workflows/main.yml:65-66).

The golden TSVs in tests/golden/*.ani.tsv.gz come from the unmodified reference CLI
(tests/golden/make_golden_tsv.py)."""
import gzip
import os
import subprocess

import numpy as np
import pytest

import make_golden as mg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
CLI = os.path.join(ROOT, "integration", "_build", "lz-ani-b200")
ALT = "--mal 15 --msl 9 --mrd 60 --mqd 60 --reg 60 --aw 20 --am 9 --ar 4".split()
LETTERS = np.frombuffer(b"ACGTNN", dtype=np.uint8)


def write_fasta(path, ss, order=None):
    ids = range(len(ss)) if order is None else order
    with open(path, "wb") as f:
        for i in ids:
            f.write(b">" + ss.names[i].encode() + b" synthetic description\n")
            s = LETTERS[ss.codes[i]].tobytes()
            for o in range(0, len(s), 70):
                f.write(s[o:o + 70] + b"\n")


def golden(name):
    return (gzip.open(os.path.join(GOLD, name + ".ani.tsv.gz")).read(),
            gzip.open(os.path.join(GOLD, name + ".ani.ids.tsv.gz")).read())


def need_cli():
    if not os.path.exists(CLI):
        pytest.skip("integration/_build/lz-ani-b200 not built (needs /root/reference at build time)")


def test_cli_has_no_cpu_path_and_fails_loudly_without_gpu(tmp_path):
    need_cli()
    syms = subprocess.run(["nm", "-C", CLI], capture_output=True, text=True).stdout
    assert "CParser::parse" not in syms and "CParser::prepare_reference" not in syms
    assert "lzani_run_rows" in syms
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present")
    except ImportError:
        pass
    ss = mg.get_set("example")
    write_fasta(tmp_path / "in.fna", ss)
    r = subprocess.run([CLI, "all2all", "--in-fasta", str(tmp_path / "in.fna"), "--out", str(tmp_path / "o.tsv"), "-V", "1"],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr
    assert not os.path.exists(tmp_path / "o.tsv") or os.path.getsize(tmp_path / "o.tsv") == 0


@pytest.mark.gpu
@pytest.mark.parametrize("case,setname,extra", [
    ("vir61_default", "vir61", []),
    ("vir61_alt_complete", "vir61", ALT + ["--out-format", "complete"]),
    ("example_default", "example", []),
    ("example_filter09", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9"]),
    # --out-filter: pairing + filter arithmetic on the device (lzani_filter_pairs), survivors only reach the host
    ("vir61_outfilter", "vir61", ["--out-filter", "ani", "0.9", "--out-filter", "qcov", "0.3"]),
    ("example_outfilter", "example", ["--out-filter", "tani", "0.85", "--out-filter", "rcov", "0.9", "--out-filter", "gani", "0.8"]),
    ("example_filter09_outfilter", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9", "--out-filter", "ani", "0.97"]),
    # submits without pairs: a filter nothing passes, a set of one sequence -- header-only output like the reference's
    ("example_filter_none", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "1.5"]),
    ("example_filter_none@2ctx", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "1.5", "--out-filter", "ani", "0.5"]),
    ("single_default", "example", []),
    ("single_default@2ctx", "example", []),
    # SURVEY 8f-3 / 8f-4 from the product binary: the reservoir keeps the loader's text and K1 packs on the device
    # (LZANI_DEVICE_INGEST); the filter's id remap + row sort + CSR are built on the device (LZANI_DEVICE_PAIRLIST)
    ("vir61_default@ingest", "vir61", []),
    ("example_default@ingest", "example", []),
    ("example_filter09@pairlist", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9"]),
    ("example_filter09_outfilter@pairlist+ingest", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9", "--out-filter", "ani", "0.97"]),
    # one library context driving two device contexts (lzani_create with n_devices = 2): rows dealt by the library
    ("vir61_default@2ctx", "vir61", []),
    ("example_filter09@2ctx", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9"]),
    # the same with the rows dealt over two contexts that read each other's results in place
    ("vir61_outfilter@2ctx", "vir61", ["--out-filter", "ani", "0.9", "--out-filter", "qcov", "0.3"]),
    ("example_filter09_outfilter@2ctx", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9", "--out-filter", "ani", "0.97"]),
    # two GPUs (skipped on a one-GPU box): rows dealt r mod 2, the filter kernel reads the peer's rows over NVLink
    ("vir61_default@2gpu", "vir61", []),
    ("vir61_outfilter@2gpu", "vir61", ["--out-filter", "ani", "0.9", "--out-filter", "qcov", "0.3"]),
    ("example_filter09_outfilter@2gpu", "example", ["--flt-kmerdb", os.path.join(GOLD, "example_fltr.txt"), "0.9", "--out-filter", "ani", "0.97"]),
])
def test_cli_tsv_byte_identical_to_reference(tmp_path, case, setname, extra):
    need_cli()
    env = dict(os.environ)
    if "@" in case and case.split("@")[1] in ("ingest", "pairlist", "pairlist+ingest"):
        case, sw = case.split("@")
        if "ingest" in sw:
            env["LZANI_DEVICE_INGEST"] = "1"
        if "pairlist" in sw:
            env["LZANI_DEVICE_PAIRLIST"] = "1"
    if case.endswith("@2ctx"):
        case = case[:-5]
        env["LZANI_DEVICES"] = "0,0"
    elif case.endswith("@2gpu"):
        import torch
        if torch.cuda.device_count() < 2:
            pytest.skip("needs two GPUs")
        case = case[:-5]
        env["LZANI_DEVICES"] = "0,1"
    ss = mg.get_set(setname)
    order = None
    if "--flt-kmerdb" in extra:
        # the filter's header lists the names in the original FASTA order (lz_matcher.cpp:43-75)
        hdr = open(os.path.join(GOLD, "example_fltr.txt")).readline().rstrip("\n").split(",")[1:]
        hdr = [h for h in hdr if h]
        order = [ss.names.index(h) for h in hdr]
    if case.startswith("single_"):
        order = [ss.names.index("NC_010807.ref")]      # the first record of example/multifasta.fna
    fa = tmp_path / "in.fna"
    write_fasta(fa, ss, order)
    out = tmp_path / "ani.tsv"
    r = subprocess.run([CLI, "all2all", "--in-fasta", str(fa), "--out", str(out), "-V", "1"] + extra,
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    want_tsv, want_ids = golden(case)
    assert open(out, "rb").read() == want_tsv
    assert open(str(out).replace("ani.tsv", "ani.ids.tsv"), "rb").read() == want_ids


@pytest.mark.gpu
@pytest.mark.parametrize("devices", ["0", "0,0"])
def test_cli_out_alignment_identical_to_reference(tmp_path, devices):
    """--out-alignment: region records from the B200 regions kernel written by the reference's own
    store_alignment; the reference's line order depends on thread interleaving, so sorted lines are
    compared.  "0,0": the library's group path (region lists of two device contexts stitched
    back into CSR order)."""
    need_cli()
    ss = mg.get_set("example")
    fa = tmp_path / "in.fna"
    write_fasta(fa, ss)
    out, aln = tmp_path / "ani.tsv", tmp_path / "ani.aln.tsv"
    r = subprocess.run([CLI, "all2all", "--in-fasta", str(fa), "--out", str(out), "--out-alignment", str(aln), "-V", "1"],
                       capture_output=True, text=True, env=dict(os.environ, LZANI_DEVICES=devices))
    assert r.returncode == 0, r.stderr
    want = gzip.open(os.path.join(GOLD, "example_default.aln.sorted.tsv.gz")).read().splitlines(keepends=True)
    got = sorted(open(aln, "rb").read().splitlines(keepends=True))
    assert got == want
    assert open(out, "rb").read() == golden("example_default")[0]
