#!/bin/bash
# usage: tools/variants.sh name "<extra nvcc flags>" ...   -- builds lz-ani_b200/variants/lib<name>.so (tuning experiments)
set -e
cd "$(dirname "$0")/.."
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
    lz-ani_b200/csrc/kernels.cu lz-ani_b200/csrc/postproc.cu lz-ani_b200/csrc/pairlist.cu lz-ani_b200/csrc/capi.cu \
    -o lz-ani_b200/variants/lib$name.so &
done
wait
