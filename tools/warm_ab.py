"""Dev tool: cold forward vs warm start (hint = the neighbours of the previous iteration) at the looped configs.
Forward only, CUDA events, best of 8.  "moved" = cloud 1 perturbed by sigma between the hint's iteration and this one."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
PEAK = 148 * 128 * 1.965e9


def best_of(fn, reps=8):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)


KIND = sys.argv[1] if len(sys.argv) > 1 else "uniform"
for name, b, n, cfg in (("C2", 32, 16384, 2), ("C3", 1, 100_000, 3), ("B=1 N=M=32768", 1, 32768, 7), ("C5-size", 1, 262_144, 5),
                        ("C4", 1, 1_000_000, 4)):
    x1, x2 = synthetic.make_pair(KIND, b, n, n, cfg)
    a, c = x1.to(dev), x2.to(dev)
    ref = F.nnd_forward_raw(a, c)
    cold = best_of(lambda: F.nnd_forward_raw(a, c), 8 if n < 500_000 else 3)
    line = f"{name:16s} cold {cold:9.4f} ms ({b * n * n * 8.0 / (cold * 1e-3) / PEAK * 100:5.1f}%)"
    for sigma in (0.0, 1e-4, 1e-3, 1e-2):
        prev = F.nnd_forward_raw(a + sigma * torch.randn_like(a), c)[3] if sigma else ref[3]
        out = F.nnd_forward_raw(a, c, hint_idx2=prev)
        assert all(torch.equal(p, q) for p, q in zip(out, ref)), "warm start changed the results"
        t = best_of(lambda: F.nnd_forward_raw(a, c, hint_idx2=prev), 8 if n < 500_000 else 3)
        line += f" | moved {sigma:g}: {t:9.4f} ms ({b * n * n * 8.0 / (t * 1e-3) / PEAK * 100:5.1f}%)"
    print(line, flush=True)
