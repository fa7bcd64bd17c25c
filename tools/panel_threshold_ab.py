"""Dev tool: persistent symmetric kernel vs panel scan (1/2/4 tiles per panel) below the 200 000-row threshold.
Forward only, CUDA events, best of 8."""
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
before = lib.nnd_set_panel_min_rows(0)
for n in (32768, 65536, 100_000, 150_000, 200_000):
    x1, x2 = synthetic.make_pair("uniform", 1, n, n, 3)
    a, c = x1.to(dev), x2.to(dev)
    ref = None
    for label, min_rows, tiles, seed in (("persistent", 1 << 40, 0, 0), ("persistent + seed 1", 1 << 40, 0, 1),
                                         ("panel x1", 1, 1, 0), ("panel x2", 1, 2, 0), ("panel x4", 1, 4, 0)):
        lib.nnd_set_panel_min_rows(min_rows)
        lib.nnd_set_panel_tiles(tiles)
        lib.nnd_set_sym_seed_tiles(seed)
        for _ in range(2):
            out = F.nnd_forward_raw(a, c)
        if ref is None:
            ref = out
        else:
            assert all(torch.equal(p, q) for p, q in zip(out, ref))
        ts = []
        for _ in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            F.nnd_forward_raw(a, c)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        best = min(ts)
        print(f"N=M={n:7d} {label:20s} forward best {best:8.4f} ms  {n * n * 8.0 / (best * 1e-3) / PEAK * 100:5.1f}% of the FP32 issue peak", flush=True)
lib.nnd_set_panel_min_rows(before)
lib.nnd_set_panel_tiles(0)
lib.nnd_set_sym_seed_tiles(1)
