"""GPU tests of the graph-captured chamfer iteration (pointcloudregistration_b200/ndp_loop.py) against the eager
pytorch3d-shaped loss + autograd (the path registration.py:236-261 takes)."""
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from pointcloudregistration_b200 import ndp_loop, synthetic  # noqa: E402
from pointcloudregistration_b200 import pytorch3d_compat as P3D  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("trunc", [1e9, 0.0005])
@pytest.mark.parametrize("use_graph", [True, False])
def test_loop_matches_eager_over_iterations(cuda_lib, device, trunc, use_graph):
    n, m = 3000, 2900
    x0, y = synthetic.make_pair("clustered", 1, n, m, 95)
    y = y.to(device)
    loop = ndp_loop.ChamferLoop(1, n, m, device, trunc=trunc, use_graph=use_graph)
    loop.set_target(y)
    xa = x0.to(device).clone()
    xb = x0.to(device).clone()
    for it in range(6):
        loss, g = loop.step(xa)
        xe = xb.clone().requires_grad_(True)
        want = P3D.compute_truncated_chamfer_distance(xe, y, trunc=trunc)
        want.backward()
        assert torch.allclose(loss, want.detach(), rtol=1e-6, atol=0), (it, float(loss), float(want))
        scale = xe.grad.abs().max()
        assert (g - xe.grad).abs().max() <= 1e-5 * scale
        xa = xa - 0.01 * g          # registration.py:259-261 (stand-in for the optimiser step)
        xb = xb - 0.01 * xe.grad
    assert loop.replays == 6
    if trunc < 1:
        assert 0 < int((loop.grad_x.abs().sum(-1) == 0).sum())  # truncated rows get no gradient


def test_autograd_function_trains_a_warp(cuda_lib, device):
    n, m = 1024, 1500
    x0, y = synthetic.make_pair("uniform", 1, n, m, 96)
    loop = ndp_loop.ChamferLoop(1, n, m, device)
    loop.set_target(y.to(device))
    shift = torch.zeros(3, device=device, requires_grad=True)  # a 3-parameter "warp": x + shift
    opt = torch.optim.Adam([shift], lr=0.01)
    first = None
    for it in range(30):
        opt.zero_grad()
        loss = ndp_loop.ChamferLoopFunction.apply(x0.to(device) + 0.2 + shift, loop)
        loss.backward()
        opt.step()
        if it % 10 == 0:  # the early-stop read, every k-th iteration only
            v = loop.loss_value()
            first = v if first is None else first
    assert loop.loss_value() < first and shift.grad is not None
