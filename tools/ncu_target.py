"""Dev tool: a short, fixed nnd forward(+backward) for ncu captures.  usage: ncu_target.py [n] [kind] [reps]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
kind = sys.argv[2] if len(sys.argv) > 2 else "uniform"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda:0")
x1, x2 = synthetic.make_pair(kind, 1, n, n, 4)
a, c = x1.to(dev), x2.to(dev)
gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(1, n, n, 4, mean=True))
for _ in range(reps):
    d1, d2, i1, i2 = F.nnd_forward_raw(a, c)
    g1, g2 = F.nnd_backward_raw(a, c, gd1, gd2, i1, i2)
torch.cuda.synchronize()
print("ok", float(d1.sum()), float(g1.abs().sum()))
