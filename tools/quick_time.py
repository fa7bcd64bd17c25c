"""Dev tool: CUDA-event timing of nnd forward / backward at a few shapes (not the bench contract)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

PEAK_LANE_OPS = 148 * 128 * 1.965e9


def timeit(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = torch.device("cuda:0")
    shapes = [(1, 2048, 2048, 200), (256, 256, 256, 100), (32, 16384, 16384, 10), (1, 100_000, 100_000, 10),
              (1, 262_144, 262_144, 5), (1, 1_000_000, 1_000_000, 3)]
    if len(sys.argv) > 1:
        shapes = [s for s in shapes if str(s[1]) in sys.argv[1:]]
    for b, n, m, iters in shapes:
        for kind in ("uniform", "clustered"):
            x1, x2 = synthetic.make_pair(kind, b, n, m, 4)
            a, c = x1.to(dev), x2.to(dev)
            out = F.nnd_forward_raw(a, c)
            gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, m, 4, mean=True))
            tf = timeit(lambda: F.nnd_forward_raw(a, c), iters)
            tb = timeit(lambda: F.nnd_backward_raw(a, c, gd1, gd2, out[2], out[3]), max(iters, 20))
            pairs = 2.0 * b * n * m
            gp = pairs / (tf * 1e-3) / 1e9
            print(f"B={b:4d} N={n:8d} M={m:8d} {kind:9s} fwd {tf:10.3f} ms  bwd {tb:8.4f} ms  "
                  f"{gp:9.1f} Gpairs/s fwd  = {gp * 8e9 / PEAK_LANE_OPS * 100:5.1f}% of FP32 lane peak; "
                  f"bwd {56.0 * b * (n + m) / (tb * 1e-3) / 1e9:7.1f} GB/s", flush=True)


if __name__ == "__main__":
    main()
