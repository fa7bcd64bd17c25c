"""Dev tool: one forward+backward at the small shapes (for an ncu launch list)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
for b, n in ((1, 2048), (1, 3000), (256, 256), (1, 16384)):
    x1, x2 = synthetic.make_pair("uniform", b, n, n, 4)
    a, c = x1.to(dev), x2.to(dev)
    gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, n, 4, mean=True))
    for _ in range(3):
        d1, d2, i1, i2 = F.nnd_forward_raw(a, c)
        F.nnd_backward_raw(a, c, gd1, gd2, i1, i2)
    torch.cuda.synchronize()
print("ok")
