"""Dev tool: atomic vs deterministic backward, CUDA-event times per call (forward indices computed once)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
for kind, b, n in (("uniform", 1, 3000), ("uniform", 256, 256), ("uniform", 1, 100_000), ("uniform", 1, 1_000_000),
                   ("clustered", 1, 1_000_000)):
    x1, x2 = (t.to(dev) for t in synthetic.make_pair(kind, b, n, n, 4))
    gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, n, 4, mean=True))
    _, _, i1, i2 = F.nnd_forward_raw(x1, x2)
    out = []
    for exact in (False, True):
        for _ in range(3):
            F.nnd_backward_raw(x1, x2, gd1, gd2, i1, i2, exact=exact)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            F.nnd_backward_raw(x1, x2, gd1, gd2, i1, i2, exact=exact)
        e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1) / 20)
    print(f"{kind:9s} B={b:4d} N=M={n:8d}   atomic {out[0]:8.4f} ms   deterministic {out[1]:8.4f} ms", flush=True)
