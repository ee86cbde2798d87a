#!/bin/bash
# usage (on a GPU box): tools/product_e2e.sh <genomes> <devices e.g. 0 or 0,1,2,3,4,5,6,7> [ref_genomes]
# Product-level end to end: the drop-in binary (the reference's own CLI / loader / writers around the B200 library)
# on a FASTA of the bench set, with the reference's own stage timers (-V 2, lz_matcher.cpp:584-614), plain and with
# --out-filter (the device-side pre-filter); optionally the unmodified reference CLI on the first
# <ref_genomes> genomes of the same FASTA (the full set is a multi-minute CPU job).
set -e
G=$1; DEV=$2; RG=${3:-0}
cd "$(dirname "$0")/.."
D=/tmp/pe2e; mkdir -p $D
python - <<PY
import sys, numpy as np
sys.path.insert(0, '.')
import bench
ss = bench.make_set($G)
L = np.frombuffer(b"ACGTNN", dtype=np.uint8)
def write(path, ids):
    with open(path, "wb") as f:
        for i in ids:
            f.write(b">" + ss.names[i].encode() + b"\n")
            s = L[ss.codes[i]].tobytes()
            f.write(b"\n".join(s[o:o + 80] for o in range(0, len(s), 80)) + b"\n")
write("$D/set.fna", range(len(ss)))
if $RG: write("$D/sub.fna", range($RG))
PY
T=$(nproc)
run() { # name, args...
  local name=$1; shift
  local t0=$(date +%s%N)
  "$@" 2> $D/$name.err > /dev/null
  local t1=$(date +%s%N)
  echo "== $name: wall $(( (t1 - t0) / 1000000 )) ms"
  grep -E "time|Time|matching|Loading|Stor|Reorder|Total|B200" $D/$name.err | tr '\r' '\n' | grep -v "^Sequences" | tail -12
}
export LZANI_DEVICES=$DEV
run "b200_plain_dev$DEV" integration/_build/lz-ani-b200 all2all --in-fasta $D/set.fna --out $D/o1.tsv -V 2 -t $T
run "b200_outfilter_dev$DEV" integration/_build/lz-ani-b200 all2all --in-fasta $D/set.fna --out $D/o2.tsv -V 2 -t $T --out-filter ani 0.7
wc -l $D/o1.tsv $D/o2.tsv
if [ "$RG" != "0" ]; then
  run "reference_cli_first$RG" oracle/_ref/lz-ani all2all --in-fasta $D/sub.fna --out $D/r.tsv -V 2 -t $T
  LZANI_DEVICES=0 run "b200_first$RG" integration/_build/lz-ani-b200 all2all --in-fasta $D/sub.fna --out $D/o3.tsv -V 2 -t $T
  cmp $D/r.tsv $D/o3.tsv && echo "TSV of the first $RG genomes: byte-identical to the reference CLI"
fi
