#!/bin/bash
# Build tuning variants of libmgb200 into build/variants/ (experiments only).
set -e
cd "$(dirname "$0")/.."
rm -rf build/variants; mkdir -p build/variants
build() { # name, extra flags
  name=$1; shift
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include "$@" \
    -shared -o build/variants/libmgb200_$name.so csrc/workspace.cu csrc/erfbin.cu csrc/allreduce.cu csrc/adam.cu csrc/linlink.cu -lcuda &
}
build w12 -DEB_TARGET_WARPS=12
build w36 -DEB_TARGET_WARPS=36 -DEB_MIN_CTAS=3
wait
ls build/variants
