"""Dev tool: C2 (B=32, N=M=16384) forward twice, for an ncu capture of the persistent symmetric kernel."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
x1, x2 = synthetic.make_pair("uniform", 32, 16384, 16384, 2)
a, c = x1.to(dev), x2.to(dev)
for _ in range(2):
    d1, d2, i1, i2 = F.nnd_forward_raw(a, c)
torch.cuda.synchronize()
print("ok", float(d1.sum()))
