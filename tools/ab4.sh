#!/bin/bash
# usage (on the GPU box): tools/ab4.sh <genomes> name...  -- config 4 (sparse phage pair list, related pairs) step time of
# lz-ani_b200/variants/lib<name>.so ("base" = the in-tree build), parity checked on a sample
G=$1; shift
for v in "$@"; do
  if [ "$v" = base ]; then unset LZANI_LIB; else export LZANI_LIB=$PWD/lz-ani_b200/variants/lib$v.so; fi
  python tools/configs45.py --config 4 --genomes $G --steps 2 --warmup 1 --no-e2e 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$v', round(d['value']), 'pairs/s  parse_ms', round(d['rank0']['parse_ms'], 1), 'index_ms', round(d['rank0']['index_ms'], 1), 'parity', d['parity']['ok'])
    elif 'rror' in l: print('$v', l.strip()[:200])
"
done
