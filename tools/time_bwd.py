"""Dev tool: CUDA-event timing of the backward only (L2-warm and after an L2 flush)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
for b, n in ((1, 1_000_000), (1, 262_144), (32, 16384), (1, 100_000), (256, 256)):
    for kind in ("uniform", "clustered"):
        x1, x2 = synthetic.make_pair(kind, b, n, n, 4)
        a, c = x1.to(dev), x2.to(dev)
        _, _, i1, i2 = F.nnd_forward_raw(a, c)
        gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, n, 4, mean=True))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        res = []
        for cold in (False, True):
            tot = 0.0
            for it in range(12):
                if cold:
                    flush.fill_(it)
                e0.record()
                F.nnd_backward_raw(a, c, gd1, gd2, i1, i2)
                e1.record()
                torch.cuda.synchronize()
                if it >= 2:
                    tot += e0.elapsed_time(e1)
            res.append(tot / 10)
        gb = 56.0 * b * 2 * n / 1e9
        print(f"B={b:4d} N=M={n:8d} {kind:9s} bwd warm {res[0]*1e3:8.1f} us ({gb/res[0]*1e3:7.1f} GB/s)  "
              f"after L2 flush {res[1]*1e3:8.1f} us ({gb/res[1]*1e3:7.1f} GB/s)", flush=True)
