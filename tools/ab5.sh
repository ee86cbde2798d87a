#!/bin/bash
# usage (on the GPU box): tools/ab5.sh <genomes> name...  -- config 5 (long-genome parameters) step time of
# lz-ani_b200/variants/lib<name>.so ("base" = the in-tree build), parity checked on a sample
G=$1; shift
for v in "$@"; do
  if [ "$v" = base ]; then unset LZANI_LIB; else export LZANI_LIB=$PWD/lz-ani_b200/variants/lib$v.so; fi
  python tools/configs45.py --config 5 --genomes $G --steps 1 --warmup 1 --no-e2e --cpu-pairs 2 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$v', round(d['value']), 'pairs/s  ms_per_step', round(d['ms_per_step'], 1), 'parity', d['parity']['ok'])
    elif 'rror' in l: print('$v', l.strip()[:200])
"
done
