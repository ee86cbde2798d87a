#!/bin/bash
# usage: tools/variants.sh name "-DMACRO=.. -DMACRO=.." [name flags ...]  -- builds lz-ani_b200/variants/lib<name>.so with extra
# nvcc flags (kernel tuning experiments; compare on the GPU box with tools/ab.sh)
set -e
cd "$(dirname "$0")/.."
mkdir -p lz-ani_b200/variants
S=lz-ani_b200/csrc
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
      $S/kernels.cu $S/postproc.cu $S/pairlist.cu $S/capi.cu $S/noalloc.cu -o lz-ani_b200/variants/lib$name.so && echo "built $name" ) &
  # at most 3 compiles at a time
  while [ "$(jobs -rp | wc -l)" -ge 3 ]; do sleep 1; done
done
wait
