#!/bin/bash
# usage (on the GPU box): tools/ab.sh <genomes> name...   -- pairs/s of lz-ani_b200/variants/lib<name>.so ("base" = the in-tree build)
G=$1; shift
for v in "$@"; do
  if [ "$v" = base ]; then unset LZANI_LIB; else export LZANI_LIB=$PWD/lz-ani_b200/variants/lib$v.so; fi
  python bench.py --genomes $G --steps 3 --warmup 2 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$v', round(d['value']), 'pairs/s  parse_ms', round(d['parse_kernel_ms_per_step'], 2), d['result_checksum'])
    elif 'rror' in l: print('$v', l.strip()[:200])
"
done
