#!/bin/bash
# usage (GPU box): tools/ab_mid.sh rounds name1 name2 ...  -- interleaved A/B of library variants at C2 / C3 (tools/sym_seed_ab.py, seed 1)
rounds="$1"; shift
for r in $(seq 1 $rounds); do for v in "$@"; do
  NND_B200_DEBUG=1 NND_B200_LIB=$PWD/tools/variants/libnnd_$v.so python tools/sym_seed_ab.py 2>/dev/null | grep "seed tiles 1" | cut -c1-100 | sed "s/^/$v: /"
done; done
