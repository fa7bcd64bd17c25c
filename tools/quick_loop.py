"""Dev tool: per-iteration time of the chamfer iteration, eager autograd path vs graph replay (ndp_loop.ChamferLoop)."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import ndp_loop, synthetic  # noqa: E402
from pointcloudregistration_b200 import pytorch3d_compat as P3D  # noqa: E402

dev = torch.device("cuda:0")
for n in (3000, 100_000):
    iters = 500
    x0, y = synthetic.make_pair("uniform", 1, n, n, 3)
    x0, y = x0.to(dev), y.to(dev)

    def eager(sync_every):
        x = x0.clone()
        for it in range(iters):
            xe = x.requires_grad_(True)
            loss = P3D.compute_truncated_chamfer_distance(xe, y, trunc=1e9)
            loss.backward()
            if it % sync_every == 0:
                loss.item()
            x = (xe - 0.01 * xe.grad).detach()

    def graphed(sync_every):
        loop = ndp_loop.ChamferLoop(1, n, n, dev)
        loop.set_target(y)
        loop.step(x0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        x = x0.clone()
        for it in range(iters):
            _, g = loop.step(x)
            if it % sync_every == 0:
                loop.loss_value()
            x = x - 0.01 * g
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / iters * 1e3

    eager(1)
    torch.cuda.synchronize()
    for k in (1, 10):
        t0 = time.perf_counter()
        eager(k)
        torch.cuda.synchronize()
        te = (time.perf_counter() - t0) / iters * 1e3
        tg = graphed(k)
        pairs = 2.0 * n * n
        print(f"N=M={n:7d} x{iters} iterations, loss read every {k:2d}: eager {te:7.3f} ms/iter ({pairs/te/1e6:8.1f} Gpairs/s)"
              f"   graph replay {tg:7.3f} ms/iter ({pairs/tg/1e6:8.1f} Gpairs/s)", flush=True)
