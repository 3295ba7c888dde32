#!/bin/bash
# Build tuning variants of libmgb200 into build/variants/ (experiments only).
set -e
cd "$(dirname "$0")/.."
rm -rf build/variants; mkdir -p build/variants
build() { # name, extra flags
  name=$1; shift
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include "$@" \
    -shared -o build/variants/libmgb200_$name.so csrc/workspace.cu csrc/erfbin.cu csrc/allreduce.cu csrc/adam.cu -lcuda &
}
build base
build et3 -DEB_ET_MAX=3
build et6 -DEB_ET_MAX=6
build et4c1 -DEB_MIN_CTAS=1 -DEB_TARGET_WARPS=12
build et6c1 -DEB_ET_MAX=6 -DEB_MIN_CTAS=1 -DEB_TARGET_WARPS=12
build et2 -DEB_ET_MAX=2
wait
ls build/variants
