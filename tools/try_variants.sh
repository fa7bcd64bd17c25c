#!/bin/bash
# usage (on the GPU box): tools/try_variants.sh outdir "<quick_time size filter>" name1 name2 ...
# For every tools/variants/libnnd_<name>.so: run the GPU parity tests against it, then time it.
out="$1"; args="$2"; shift 2
mkdir -p "$out"
for v in "$@"; do
  lib=$PWD/tools/variants/libnnd_$v.so
  echo "=== variant $v" | tee -a "$out/summary.txt"
  NND_B200_DEBUG=1 NND_B200_LIB=$lib timeout 600 python -m pytest tests/test_nnd_gpu.py -x -q -m gpu 2>&1 | tail -2 | tee -a "$out/summary.txt"
  NND_B200_DEBUG=1 NND_B200_LIB=$lib timeout 600 python tools/quick_time.py $args 2>&1 | grep -v clustered | tee -a "$out/summary.txt"
done
