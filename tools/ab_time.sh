#!/bin/bash
# usage (GPU box): tools/ab_time.sh "<quick_time size filter>" rounds name1 name2 ...   -- interleaved A/B timing
args="$1"; rounds="$2"; shift 2
for r in $(seq 1 $rounds); do for v in "$@"; do
  NND_B200_DEBUG=1 NND_B200_LIB=$PWD/tools/variants/libnnd_$v.so python tools/quick_time.py $args 2>&1 | grep uniform | cut -c1-75 | sed "s/^/$v: /"
done; done
