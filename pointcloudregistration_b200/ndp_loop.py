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

Forward and backward are the kernels behind pytorch3d_compat.compute_truncated_chamfer_distance; the mask, the
mean/mean reduction and d loss / d dist are one fused launch (nnd_chamfer_reduce_cuda_f32: float64 partial sums in a
fixed order, so the loss is deterministic and agrees with the eager torch reduction to fp32 rounding).
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import _cabi
from .functional import _stream_ptr, chamfer_reduce_workspace


class ChamferLoop:
    def __init__(self, b: int, n: int, m: int, device, trunc: float = 1e9, use_graph: bool = True):
        self.b, self.n, self.m, self.trunc = b, n, m, float(trunc)
        self.device = dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("ChamferLoop needs a CUDA device: pointcloudregistration_b200 has no CPU path")
        self.lib = _cabi.load()
        f32, i32 = dict(dtype=torch.float32, device=dev), dict(dtype=torch.int32, device=dev)
        self.x = torch.zeros(b, n, 3, **f32)
        self.y = torch.zeros(b, m, 3, **f32)
        self.loss = torch.zeros((), **f32)
        self.grad_x = torch.zeros_like(self.x)
        self.grad_y = torch.zeros_like(self.y)
        self.dist1, self.dist2 = torch.zeros(b, n, **f32), torch.zeros(b, m, **f32)
        self.idx1, self.idx2 = torch.zeros(b, n, **i32), torch.zeros(b, m, **i32)
        self._gd1, self._gd2 = torch.zeros(b, n, **f32), torch.zeros(b, m, **f32)
        self._ws = torch.empty(max(self.lib.nnd_workspace_bytes(b, n, m), 256), dtype=torch.uint8, device=dev)
        self._rws = chamfer_reduce_workspace(dev)
        self.graph = None
        self.use_graph = use_graph
        self.replays = 0

    # what one iteration computes (also the eager path when use_graph is False): 3-4 forward launches, the fused
    # mask + mean/mean reduction + d loss/d dist, 2 memsets + the gradient kernel -- all over static buffers
    def _body(self):
        lib, st = self.lib, _stream_ptr(self.device)
        b, n, m = self.b, self.n, self.m
        rc = lib.nnd_forward_cuda_f32(self.x.data_ptr(), self.y.data_ptr(), b, n, m, self.dist1.data_ptr(),
                                      self.dist2.data_ptr(), self.idx1.data_ptr(), self.idx2.data_ptr(),
                                      self._ws.data_ptr(), self._ws.numel(), st)
        _cabi.check(rc, "nnd_forward_cuda_f32")
        # point_reduction = mean (divide by the cloud length), batch_reduction = mean, both directions summed
        s1, s2 = 1.0 / (max(n, 1) * max(b, 1)), 1.0 / (max(m, 1) * max(b, 1))
        rc = lib.nnd_chamfer_reduce_cuda_f32(self.dist1.data_ptr(), self.dist2.data_ptr(), b, n, m, self.trunc, s1, s2,
                                             self.loss.data_ptr(), self._gd1.data_ptr(), self._gd2.data_ptr(),
                                             self._rws.data_ptr(), self._rws.numel(), st)
        _cabi.check(rc, "nnd_chamfer_reduce_cuda_f32")
        rc = lib.nnd_backward_cuda_f32(self.x.data_ptr(), self.y.data_ptr(), self._gd1.data_ptr(), self._gd2.data_ptr(),
                                       self.idx1.data_ptr(), self.idx2.data_ptr(), b, n, m, self.grad_x.data_ptr(),
                                       self.grad_y.data_ptr(), st)
        _cabi.check(rc, "nnd_backward_cuda_f32")

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
