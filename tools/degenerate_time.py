"""Dev tool: forward time on degenerate clouds (all points identical / few distinct points)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F  # noqa: E402

dev = torch.device("cuda:0")
n = 262144
for name, gen in (("all identical", lambda: torch.full((1, n, 3), 0.25)),
                  ("16 distinct points", lambda: torch.randint(0, 16, (1, n, 1)).float().expand(1, n, 3).contiguous() / 16),
                  ("uniform", lambda: torch.rand(1, n, 3))):
    torch.manual_seed(0)
    a, c = gen().to(dev), gen().to(dev)
    F.nnd_forward_raw(a, c)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    d1, d2, i1, i2 = F.nnd_forward_raw(a, c)
    e1.record()
    torch.cuda.synchronize()
    print(f"N=M={n} {name:20s} fwd {e0.elapsed_time(e1):9.3f} ms   idx1 max {int(i1.max())} idx2 max {int(i2.max())}", flush=True)
