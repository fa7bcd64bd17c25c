"""Dev tool: a short, fixed run for the round-2 ncu captures: the single-launch small-cloud kernels at the DIP shape and
C1, and the persistent symmetric kernel at C2 (forward + backward each, twice)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
for b, n, cfg in ((256, 256, 6), (1, 2048, 1), (32, 16384, 2)):
    x1, x2 = synthetic.make_pair("uniform", b, n, n, cfg)
    a, c = x1.to(dev), x2.to(dev)
    gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, n, cfg, mean=True))
    for _ in range(2):
        d1, d2, i1, i2 = F.nnd_forward_raw(a, c)
        g1, g2 = F.nnd_backward_raw(a, c, gd1, gd2, i1, i2)
    torch.cuda.synchronize()
    print("ok", b, n, float(d1.sum()), float(g1.abs().sum()))
