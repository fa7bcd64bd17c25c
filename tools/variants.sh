#!/bin/bash
# Dev tool: build experimental variants of libnnd_b200.so into tools/variants/ (git-ignored *.so).
# usage: tools/variants.sh "NAME:-DFLAG=1 -DOTHER=2" ...
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/variants
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -shared -Iinclude $flags \
    -o tools/variants/libnnd_$name.so pointcloudregistration_b200/csrc/nnd_kernels.cu 2>&1 | grep -v -E "warning|fmin3|\^|^$|Remark" || true
  echo "built tools/variants/libnnd_$name.so ($flags)"
done
