"""Repeated chamfer on one fixed-size pair without per-iteration host work.

The reference's non-rigid stage (c2p-net/deformationpyramid/model/registration.py:208-262) evaluates
compute_truncated_chamfer_distance(s_warp, t_sample, trunc=1e9) (registration.py:236), calls loss.item() for the
early-stop test (registration.py:250-256: a device->host sync every iteration) and back-propagates, up to 500 times
per pyramid level on clouds whose sizes never change.  At the real cloud size (~3 k points) the kernels take tens of
microseconds and the iteration is dominated by Python, allocator and launch overhead.

ChamferLoop captures  nnd forward -> truncation mask + mean/mean reduction -> nnd backward  ONCE into a CUDA graph
over static buffers and replays it: one graph launch per iteration, no allocation, no host synchronisation.  The loss
stays on the device; read it every k-th iteration (`loss_value()`), which is what makes the early-stop test cheap.
`ChamferLoopFunction` plugs the replay into autograd so the warp network in front of it trains as before.

Numerics are those of pytorch3d_compat.compute_truncated_chamfer_distance (same kernels, same reduction order).
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from .functional import nnd_backward_raw, nnd_forward_raw


class ChamferLoop:
    def __init__(self, b: int, n: int, m: int, device, trunc: float = 1e9, use_graph: bool = True):
        self.b, self.n, self.m, self.trunc = b, n, m, float(trunc)
        self.device = torch.device(device)
        self.x = torch.zeros(b, n, 3, device=self.device)
        self.y = torch.zeros(b, m, 3, device=self.device)
        self.loss = torch.zeros((), device=self.device)
        self.grad_x = torch.zeros_like(self.x)
        self.grad_y = torch.zeros_like(self.y)
        self.idx1 = torch.zeros(b, n, dtype=torch.int32, device=self.device)
        self.idx2 = torch.zeros(b, m, dtype=torch.int32, device=self.device)
        self.graph = None
        self.use_graph = use_graph
        self.replays = 0

    # what one iteration computes (also the eager path when use_graph is False)
    def _body(self):
        d1, d2, i1, i2 = nnd_forward_raw(self.x, self.y)
        k1, k2 = d1 < self.trunc, d2 < self.trunc  # loss.py:151-154: distances >= trunc do not count
        z1, z2 = torch.where(k1, d1, 0.0), torch.where(k2, d2, 0.0)
        # point_reduction = mean (divide by the cloud length), batch_reduction = mean, both directions summed
        self.loss.copy_((z1.sum(1) / self.n).sum() / self.b + (z2.sum(1) / self.m).sum() / self.b)
        gd1 = k1.to(torch.float32) / (self.n * self.b)
        gd2 = k2.to(torch.float32) / (self.m * self.b)
        g1, g2 = nnd_backward_raw(self.x, self.y, gd1, gd2, i1, i2)
        self.grad_x.copy_(g1)
        self.grad_y.copy_(g2)
        self.idx1.copy_(i1)
        self.idx2.copy_(i2)

    def _capture(self):
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):  # warm-up outside capture (lazy module loads, func attributes)
            for _ in range(2):
                self._body()
        torch.cuda.current_stream(self.device).wait_stream(s)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body()

    def set_target(self, y: torch.Tensor) -> None:
        self.y.copy_(y.detach().reshape(self.b, self.m, 3))

    def step(self, x: torch.Tensor):
        """loss (0-dim device tensor, valid until the next step) and d loss / d x [b,n,3]; no host sync."""
        self.x.copy_(x.detach().reshape(self.b, self.n, 3))
        with torch.cuda.device(self.device):
            if not self.use_graph:
                self._body()
            else:
                if self.graph is None:
                    self._capture()
                self.graph.replay()
        self.replays += 1
        return self.loss, self.grad_x

    def loss_value(self) -> float:
        """Device -> host read of the current loss (synchronises): call it every k-th iteration only."""
        return float(self.loss.item())


class ChamferLoopFunction(Function):
    """loss = ChamferLoopFunction.apply(x, loop): differentiable w.r.t. x (the warped source), target fixed."""

    @staticmethod
    def forward(ctx, x, loop: ChamferLoop):
        loss, grad_x = loop.step(x)
        ctx.save_for_backward(grad_x.clone())
        ctx.shape = x.shape
        return loss.clone()

    @staticmethod
    def backward(ctx, g):
        (grad_x,) = ctx.saved_tensors
        return (grad_x * g).reshape(ctx.shape), None
