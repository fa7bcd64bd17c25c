"""Dev tool: automatic column-filter hints (csrc/nnd_hint.cuh) off / panel scans only / persistent kernel too.
Forward only, CUDA events, best of N."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import _cabi  # noqa: E402
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

lib = _cabi.load()
dev = torch.device("cuda:0")
PEAK = 148 * 128 * 1.965e9
keep = lib.nnd_set_auto_hint(-1)
for name, kind, b, n, cfg in (("C2", "uniform", 32, 16384, 2), ("C3", "uniform", 1, 100_000, 3), ("150k", "uniform", 1, 150_000, 3),
                              ("C5 lattice", "lattice", 1, 262_144, 5), ("C5 clustered", "clustered", 1, 262_144, 5),
                              ("C5 crosscopy", "crosscopy", 1, 262_144, 5), ("262144 uniform", "uniform", 1, 262_144, 5),
                              ("C4 uniform", "uniform", 1, 1_000_000, 4), ("C4 clustered", "clustered", 1, 1_000_000, 4)):
    x1, x2 = synthetic.make_pair(kind, b, n, n, cfg)
    a, c = x1.to(dev), x2.to(dev)
    ref, line = None, f"{name:16s}"
    for mode in (0, 1, 2):
        lib.nnd_set_auto_hint(mode)
        out = F.nnd_forward_raw(a, c)
        if ref is None:
            ref = out
        else:
            assert all(torch.equal(p, q) for p, q in zip(out, ref)), "hints changed the results"
        ts = []
        for _ in range(6 if n < 500_000 else 3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            F.nnd_forward_raw(a, c)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = min(ts)
        line += f" | mode {mode}: {t:9.4f} ms ({b * n * n * 8.0 / (t * 1e-3) / PEAK * 100:5.1f}%)"
    print(line, flush=True)
lib.nnd_set_auto_hint(keep)
