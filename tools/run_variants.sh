#!/bin/bash
# usage (on the GPU box): tools/run_variants.sh "<quick_time args>" name1 name2 ...
args="$1"; shift
for v in "$@"; do
  echo "=== variant $v"
  NND_B200_LIB=$PWD/tools/variants/libnnd_$v.so python tools/quick_time.py $args 2>&1 | grep -v clustered
done
