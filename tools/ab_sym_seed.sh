#!/bin/bash
# Round-2 A/B (GPU box): persistent symmetric scan with and without the filter seed (NND_B200_SYM_SEED = sample tiles),
# at C2 (B=32, N=M=16384) and C3 (N=M=100000).  Results are bit-identical either way (tests/test_nnd_gpu.py covers the
# seeded filter through the panel path); this only measures time.  usage: bash tools/ab_sym_seed.sh [rounds]
rounds=${1:-2}
for r in $(seq 1 $rounds); do for seed in 0 1 2 4; do
  NND_B200_SYM_SEED=$seed python tools/quick_time.py 16384 100000 2>&1 | grep uniform | cut -c1-75 | sed "s/^/seed=$seed: /"
done; done
