#!/bin/bash
# Build tuning variants of libmgb200 into build/variants/ (experiments only).
# usage: tools/build_variants.sh name1 "flags1" name2 "flags2" ...
set -e
cd "$(dirname "$0")/.."
mkdir -p build/variants
build() { # name, extra flags
  name=$1; shift
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include $@ \
    -shared -o build/variants/libmgb200_$name.so csrc/workspace.cu csrc/erfbin.cu csrc/allreduce.cu csrc/adam.cu csrc/linlink.cu csrc/group.cu csrc/nvls.cu &
}
while [ $# -ge 2 ]; do build "$1" $2; shift 2; done
wait
ls build/variants
