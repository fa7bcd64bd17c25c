"""Dev tool: A/B of the column-filter seed in front of the PERSISTENT symmetric kernel (nnd_set_sym_seed_tiles) at the
mid-size configs C2 / C3.  Forward only, CUDA events, best and mean of 10."""
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
for name, b, n, cfg in (("C2", 32, 16384, 2), ("C3", 1, 100_000, 3), ("B=1 N=M=32768", 1, 32768, 7), ("B=8 N=M=16384", 8, 16384, 8)):
    x1, x2 = synthetic.make_pair("uniform", b, n, n, cfg)
    a, c = x1.to(dev), x2.to(dev)
    ref = None
    for tiles in (0, 1, 2, 4, 8):
        lib.nnd_set_sym_seed_tiles(tiles)
        for _ in range(3):
            out = F.nnd_forward_raw(a, c)
        if ref is None:
            ref = out
        else:
            assert all(torch.equal(p, q) for p, q in zip(out, ref)), "seeded scan changed the results"
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            F.nnd_forward_raw(a, c)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        best, mean = min(ts), sum(ts) / len(ts)
        print(f"{name:16s} seed tiles {tiles}: forward best {best:8.4f} ms  mean {mean:8.4f} ms  "
              f"{b * n * n * 8 / (best * 1e-3) / PEAK * 100:5.1f}% of the FP32 issue peak (evaluations executed)", flush=True)
    lib.nnd_set_sym_seed_tiles(1)
