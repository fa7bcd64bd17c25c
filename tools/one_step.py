"""Dev tool: one forward + backward of a bench workload (for ncu launch lists).  usage: one_step.py c2|c3|c5|c4 [kind]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from pointcloudregistration_b200 import functional as F, synthetic  # noqa: E402

w = dict(bench.WORKLOADS[sys.argv[1]])
if len(sys.argv) > 2:
    w["kind"] = sys.argv[2]
dev = torch.device("cuda:0")
x, y = synthetic.make_pair(w["kind"], w["b"], w["n"], w["m"], w["config"])
g1, g2 = synthetic.grad_pair(w["b"], w["n"], w["m"], w["config"], mean=True)
x, y, g1, g2 = x.to(dev), y.to(dev), g1.to(dev), g2.to(dev)
for _ in range(2):
    d1, d2, i1, i2 = F.nnd_forward_raw(x, y)
    F.nnd_backward_raw(x, y, g1, g2, i1, i2)
torch.cuda.synchronize()
