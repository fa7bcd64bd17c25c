"""Dev tool: where the end-to-end step (host buffers in, host buffers out) spends its time."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch_nndistance as NND  # noqa: E402
from pointcloudregistration_b200 import synthetic  # noqa: E402

dev = torch.device("cuda:0")
n = 1_000_000
x1h, x2h = synthetic.make_pair("uniform", 1, n, n, 4)
x1h, x2h = x1h.pin_memory(), x2h.pin_memory()
g1h = torch.empty(1, n, 3).pin_memory()
g2h = torch.empty(1, n, 3).pin_memory()


def t(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


a = x1h.to(dev)
c = x2h.to(dev)
print(f"H2D 2 x 12 MB pinned      {t(lambda: (x1h.to(dev, non_blocking=True), x2h.to(dev, non_blocking=True))):8.3f} ms")
ga = torch.rand(1, n, 3, device=dev)
print(f"D2H 2 x 12 MB pinned      {t(lambda: (g1h.copy_(ga, non_blocking=True), g2h.copy_(ga, non_blocking=True))):8.3f} ms")
print(f"nnd forward (no autograd) {t(lambda: NND.nnd(a, c)):8.3f} ms")


def fb():
    p = a.detach().requires_grad_(True)
    q = c.detach().requires_grad_(True)
    d1, d2 = NND.nnd(p, q)
    (d1.mean() + d2.mean()).backward()
    return p.grad


print(f"nnd forward + backward    {t(fb):8.3f} ms")


def full():
    p = x1h.to(dev, non_blocking=True).requires_grad_(True)
    q = x2h.to(dev, non_blocking=True).requires_grad_(True)
    d1, d2 = NND.nnd(p, q)
    loss = d1.mean() + d2.mean()
    loss.backward()
    g1h.copy_(p.grad, non_blocking=True)
    g2h.copy_(q.grad, non_blocking=True)
    return loss.item()


print(f"full e2e step             {t(full):8.3f} ms")
