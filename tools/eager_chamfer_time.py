"""Dev tool: eager pytorch3d-shaped chamfer_distance + backward per call at the callers' real shapes."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import synthetic  # noqa: E402
from pointcloudregistration_b200 import pytorch3d_compat as P3D  # noqa: E402

dev = torch.device("cuda:0")
for b, n, m in ((256, 256, 256), (1, 3000, 3000), (1, 100_000, 100_000)):
    x, y = synthetic.make_pair("uniform", b, n, m, 3)
    x, y = x.to(dev).requires_grad_(True), y.to(dev)
    for fn, name in ((lambda: P3D.chamfer_distance(x, y)[0], "chamfer_distance"),
                     (lambda: P3D.compute_truncated_chamfer_distance(x, y, trunc=1e9), "truncated"),
                     (lambda: P3D.chamfer_distance(x, y, x_lengths=torch.full((b,), n, device=dev))[0], "general path")):
        for _ in range(5):
            x.grad = None
            fn().backward()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        iters = 200
        for _ in range(iters):
            x.grad = None
            fn().backward()
        torch.cuda.synchronize()
        print(f"B={b} N={n} M={m} {name:18s} fwd+bwd {(time.perf_counter()-t0)/iters*1e3:7.3f} ms/call", flush=True)
