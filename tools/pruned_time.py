"""Dev tool (GPU box): exact pruned search (nnd_forward_pruned_cuda_f32) against the brute-force forward, same bits.
Forward only, CUDA events, best of N."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")


def best_of(fn, reps):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)


shapes = [("C2 uniform", "uniform", 32, 16384, 2), ("C3 uniform", "uniform", 1, 100_000, 3),
          ("C5 lattice", "lattice", 1, 262_144, 5), ("C5 duplicates", "duplicates", 1, 262_144, 5),
          ("C5 identical", "identical", 1, 262_144, 5), ("C5 crosscopy", "crosscopy", 1, 262_144, 5),
          ("C5 clustered", "clustered", 1, 262_144, 5), ("C4 uniform", "uniform", 1, 1_000_000, 4),
          ("C4 clustered", "clustered", 1, 1_000_000, 4), ("16k lone", "uniform", 1, 16384, 2),
          ("DIP-like 256x2048", "uniform", 256, 2048, 6)]
only = sys.argv[1:]
for name, kind, b, n, cfg in shapes:
    if only and not any(o in name for o in only):
        continue
    x1, x2 = synthetic.make_pair(kind, b, n, n, cfg)
    a, c = x1.to(dev), x2.to(dev)
    want = F.nnd_forward_raw(a, c)
    got = F.nnd_forward_pruned_raw(a, c)
    same = all(torch.equal(g, w) for g, w in zip(got, want))
    big = n >= 500_000
    tb = best_of(lambda: F.nnd_forward_raw(a, c), 2 if big else 5)
    tp = best_of(lambda: F.nnd_forward_pruned_raw(a, c), 5)
    print(f"{name:20s} B={b:3d} N=M={n:8d}: brute {tb:9.3f} ms | pruned {tp:9.3f} ms | x{tb / tp:7.1f} | "
          f"{'bit-identical' if same else 'DIFFERENT'}", flush=True)

if not only or "10M" in only:  # beyond the brute-force scan's reach (~25 s per forward): pruned only
    for kind in ("uniform", "clustered"):
        n = 10_000_000
        x1, x2 = synthetic.make_pair(kind, 1, n, n, 4)
        a, c = x1.to(dev), x2.to(dev)
        tp = best_of(lambda: F.nnd_forward_pruned_raw(a, c), 3)
        print(f"10M {kind:12s}     B=  1 N=M={n:8d}: brute (extrapolated) {243.0 * 100:9.0f} ms | pruned {tp:9.3f} ms", flush=True)
