#!/usr/bin/env python
"""profiles/traffic.json from an ncu launch list of one bench step (the `--metrics gpu__time_duration.sum,...,
dram__bytes_read.sum,dram__bytes_write.sum --clock-control none` pass of B200_PROFILING.md):

    ncu --metrics ... --csv --log-file gpurun_out/X_launches.csv python bench.py --genomes 2000 --steps 1 --warmup 0 --no-parity --no-e2e
    python tools/traffic.py gpurun_out/X_launches.csv 2000 > profiles/traffic.json

DRAM bytes of the parse stage of ONE step (every staged kernel, index build excluded), per kernel and in total, stamped
with the digest of the kernel sources (bench.csrc_digest): bench.py reports roofline.traffic only while that digest
matches the sources it runs."""
import collections
import csv
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

path, genomes = sys.argv[1], int(sys.argv[2])
rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
hdr = rows[0]
ik, im, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: collections.defaultdict(float))
for r in rows[1:]:
    k = r[ik].split("(")[0].replace("void ", "").replace("lzani::", "").split("<")[0]
    if k.startswith("cub::"):
        k = "DeviceScan" if "Scan" in k else "cub"
    agg[k][r[im]] += float(r[iv].replace(",", ""))
skip = ("k_pack_sequences", "k_build_index")
rd = sum(m["dram__bytes_read.sum"] for k, m in agg.items() if k not in skip)
wr = sum(m["dram__bytes_write.sum"] for k, m in agg.items() if k not in skip)
out = {"kernel": "parse stage of one step (all staged kernels: " + ", ".join(k for k in agg if k not in skip) + ")",
       "genomes": genomes, "n_gpus": 1, "csrc_sha1": bench.csrc_digest(),
       "dram_bytes_read": int(rd), "dram_bytes_write": int(wr), "dram_bytes_per_launch": int(rd + wr),
       "dram_bytes_by_kernel": {k: int(m["dram__bytes_read.sum"] + m["dram__bytes_write.sum"]) for k, m in agg.items() if k not in skip},
       "launch_ns_under_ncu": int(sum(m["gpu__time_duration.sum"] for k, m in agg.items() if k not in skip)),
       "source": f"{os.path.basename(path)} (ncu launch list of python bench.py --genomes {genomes} --steps 1 --warmup 0 --no-parity --no-e2e)"}
print(json.dumps(out, indent=1))
