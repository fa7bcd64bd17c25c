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
from .functional import _pruned_route, _stream_ptr, chamfer_reduce_workspace


class ChamferLoop:
    def __init__(self, b: int, n: int, m: int, device, trunc: float = 1e9, use_graph: bool = True, warm: bool = True,
                 search: str = None, resort_every: int = 32):
        """search: 'brute' (the reference's scan), 'pruned' (the exact pruned search, same bits -- DESIGN.md 4.7), or
        None = functional.search_mode() with its size thresholds.
        resort_every (pruned search): the target cloud is prepared once per set_target(); the moving cloud keeps its curve
        order between steps and is re-sorted every resort_every-th step (nnd_forward_pruned_reuse_cuda_f32: the results
        never depend on it, a stale order only prunes less).  0 = sort both clouds in every step."""
        self.b, self.n, self.m, self.trunc = b, n, m, float(trunc)
        self.warm = bool(warm)  # hand the previous iteration's idx2 to the forward as a hint (results never depend on it)
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
        if search not in (None, "brute", "pruned"):
            raise ValueError("search must be None, 'brute' or 'pruned'")
        self.pruned = search == "pruned" or (search is None and _pruned_route(b, n, m))
        nbytes = self.lib.nnd_pruned_workspace_bytes(b, n, m) if self.pruned else self.lib.nnd_workspace_bytes(b, n, m)
        if self.pruned and (nbytes == 0 or b > 65535):
            raise ValueError("the pruned search needs b <= 65535 and b*n, b*m < 2^31")
        self._ws = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
        self._rws = chamfer_reduce_workspace(dev)
        self.graph = None   # brute-force body / the pruned body with reuse = 0
        self._graphs = {}   # pruned search: one captured body per `reuse` value (0, 2, 3)
        self.use_graph = use_graph
        self.replays = 0
        self.resort_every = max(int(resort_every), 0)
        self._target_ready = False  # the workspace holds the current target's tiles
        self._since_sort = 0

    # what one iteration computes (also the eager path when use_graph is False): 3-4 forward launches, the fused
    # mask + mean/mean reduction + d loss/d dist, 2 memsets + the gradient kernel -- all over static buffers
    def _body(self, reuse: int = 0):
        lib, st = self.lib, _stream_ptr(self.device)
        b, n, m = self.b, self.n, self.m
        # warm start: the loop's own idx2 of the previous iteration is the hint (any content is safe, results do not
        # depend on it; the first iteration sees zeros = "try row 0")
        self._forward(st, reuse)
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

    def _forward(self, st, reuse: int = 0) -> None:
        """nnd forward over the static buffers on raw stream `st` (capturable: no allocation, no host read)."""
        lib, b, n, m = self.lib, self.b, self.n, self.m
        if self.pruned:
            rc = lib.nnd_forward_pruned_reuse_cuda_f32(self.x.data_ptr(), self.y.data_ptr(), b, n, m, self.dist1.data_ptr(),
                                                       self.dist2.data_ptr(), self.idx1.data_ptr(), self.idx2.data_ptr(),
                                                       self._ws.data_ptr(), self._ws.numel(), int(reuse), st)
            _cabi.check(rc, "nnd_forward_pruned_reuse_cuda_f32")
            return
        rc = lib.nnd_forward_warm_cuda_f32(self.x.data_ptr(), self.y.data_ptr(), b, n, m, self.dist1.data_ptr(),
                                           self.dist2.data_ptr(), self.idx1.data_ptr(), self.idx2.data_ptr(),
                                           self.idx2.data_ptr() if self.warm else None, self._ws.data_ptr(),
                                           self._ws.numel(), st)
        _cabi.check(rc, "nnd_forward_warm_cuda_f32")

    def _capture(self, reuse: int = 0):
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):  # warm-up outside capture (lazy module loads, func attributes)
            for _ in range(2):
                self._body(reuse)
        torch.cuda.current_stream(self.device).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._body(reuse)
        self._graphs[reuse] = g
        if reuse == 0:
            self.graph = g

    def _next_reuse(self) -> int:
        """Which parts of the previous step's preparation this step may reuse (pruned search only)."""
        if not self.pruned or self.resort_every == 0 or not self._target_ready:
            r = 0                      # everything from scratch: the first step after set_target()
        elif self._since_sort + 1 >= self.resort_every:
            r = 2                      # target tiles kept, the moving cloud re-sorted
        else:
            r = 3                      # ... and its order kept too
        self._since_sort = 0 if r != 3 else self._since_sort + 1
        self._target_ready = True
        return r

    def set_target(self, y: torch.Tensor) -> None:
        self.y.copy_(y.detach().reshape(self.b, self.m, 3), non_blocking=True)
        self._target_ready = False

    def step(self, x: torch.Tensor):
        """loss (0-dim device tensor, valid until the next step) and d loss / d x [b,n,3]; no host sync."""
        self.x.copy_(x.detach().reshape(self.b, self.n, 3), non_blocking=True)  # stream-ordered, also from pinned host memory
        reuse = self._next_reuse()
        with torch.cuda.device(self.device):
            if not self.use_graph:
                self._body(reuse)
            else:
                if reuse not in self._graphs:
                    self._capture(reuse)
                self._graphs[reuse].replay()
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


class _StaticChamfer(Function):
    """loss = truncated chamfer(x_warped, target) over STATIC buffers through the C ABI -- the piece of
    GraphedIteration that sits between the warp network's forward and its backward.  Everything it launches is
    capturable (no allocation, no host read): nnd forward, the fused mean/mean epilogue (which also leaves
    d loss / d dist), and in backward the nnd backward (deterministic variant when enabled)."""

    @staticmethod
    def forward(ctx, x, st):
        lib, dev = st.lib, st.device
        b, n, m = st.b, st.n, st.m
        st.x.copy_(x.reshape(b, n, 3))
        sp = _stream_ptr(dev)
        st._forward(sp)  # warm start / pruned search as the loop was built
        s1, s2 = 1.0 / (max(n, 1) * max(b, 1)), 1.0 / (max(m, 1) * max(b, 1))
        rc = lib.nnd_chamfer_reduce_cuda_f32(st.dist1.data_ptr(), st.dist2.data_ptr(), b, n, m, st.trunc, s1, s2,
                                             st.loss.data_ptr(), st._gd1.data_ptr(), st._gd2.data_ptr(),
                                             st._rws.data_ptr(), st._rws.numel(), sp)
        _cabi.check(rc, "nnd_chamfer_reduce_cuda_f32")
        ctx.st, ctx.shape = st, x.shape
        return st.loss.clone()

    @staticmethod
    def backward(ctx, g):
        st = ctx.st
        lib, dev = st.lib, st.device
        b, n, m = st.b, st.n, st.m
        sp = _stream_ptr(dev)
        if st.exact:
            rc = lib.nnd_backward_exact_cuda_f32(st.x.data_ptr(), st.y.data_ptr(), st._gd1.data_ptr(), st._gd2.data_ptr(),
                                                 st.idx1.data_ptr(), st.idx2.data_ptr(), b, n, m, st.grad_x.data_ptr(),
                                                 st.grad_y.data_ptr(), st._ews.data_ptr(), st._ews.numel(), sp)
            _cabi.check(rc, "nnd_backward_exact_cuda_f32")
        else:
            rc = lib.nnd_backward_cuda_f32(st.x.data_ptr(), st.y.data_ptr(), st._gd1.data_ptr(), st._gd2.data_ptr(),
                                           st.idx1.data_ptr(), st.idx2.data_ptr(), b, n, m, st.grad_x.data_ptr(),
                                           st.grad_y.data_ptr(), sp)
            _cabi.check(rc, "nnd_backward_cuda_f32")
        return (st.grad_x * g).reshape(ctx.shape), None


class GraphedIteration:
    """The WHOLE iteration of the reference's non-rigid stage (registration.py:231-261) as one CUDA graph:

        s_warped = warp(source)                                           registration.py:232  (any nn.Module / callable)
        loss     = compute_truncated_chamfer_distance(s_warped, target)   registration.py:236  (+ extra_loss(), e.g. the
                                                                          w_reg BCE term of :241-245)
        optimizer.zero_grad(); loss.backward(); optimizer.step()          registration.py:258-260

    The reference reads loss.item() twice per iteration for its early-stop test (registration.py:250-256): a host
    synchronisation per iteration, which at the real ~3 k-point size dominates the iteration.  Here the loss stays on the
    device; run() applies the SAME stop rules (loss < 1e-4; break_counter of relative changes below the threshold) but
    reads the loss every `read_every`-th iteration only.

    warp's parameters are optimised in place by `optimizer` (built by the caller, e.g. torch.optim.Adam(params,
    lr=..., capturable=True) as torch requires for graph capture).  `use_graph=False` runs the identical sequence
    eagerly (the comparison path of the tests).  exact=True takes the deterministic backward (bit-reproducible runs)."""

    def __init__(self, warp, optimizer, source: torch.Tensor, target: torch.Tensor, trunc: float = 1e9,
                 extra_loss=None, use_graph: bool = True, exact: bool = False, search: str = None):
        dev = source.device
        if dev.type != "cuda":
            raise RuntimeError("GraphedIteration needs CUDA tensors: pointcloudregistration_b200 has no CPU path")
        n, m = source.shape[-2], target.shape[-2]
        self.st = st = ChamferLoop(1, n, m, dev, trunc=trunc, use_graph=False, search=search)
        st.exact = bool(exact)
        if st.exact:
            nbytes = st.lib.nnd_backward_exact_workspace_bytes(1, n, m)
            st._ews = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
        st.set_target(target)
        self.warp, self.optimizer, self.extra_loss = warp, optimizer, extra_loss
        self.source = source.detach().reshape(n, 3).clone()
        self.warped = torch.zeros_like(self.source)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.use_graph, self.graph = use_graph, None
        self.iterations = 0

    def _body(self):
        self.optimizer.zero_grad(set_to_none=False)
        y = self.warp(self.source)
        loss = _StaticChamfer.apply(y, self.st)
        if self.extra_loss is not None:
            loss = loss + self.extra_loss()
        loss.backward()
        self.optimizer.step()
        self.warped.copy_(y.detach())
        self.loss.copy_(loss.detach())

    def _capture(self):
        dev = self.source.device
        # torch's recipe for whole-step capture: a few eager iterations on a side stream first (lazy initialisation of
        # cuBLAS handles, optimizer state, our func attributes), with parameters and optimizer state restored afterwards
        # so that the captured run starts from the caller's state
        params = [p for g in self.optimizer.param_groups for p in g["params"]]
        saved = [p.detach().clone() for p in params]
        saved_opt = self.optimizer.state_dict()
        import copy

        saved_opt = copy.deepcopy(saved_opt)
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            for g in self.optimizer.param_groups:
                for p in g["params"]:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
            for _ in range(3):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(s)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body()
        with torch.no_grad():  # the warm-up and the capture pass moved the parameters: put the caller's state back
            for p, v in zip(params, saved):
                p.copy_(v)
            live = self.optimizer.state_dict()["state"]
            for k, stt in saved_opt["state"].items():
                for name, v in stt.items():
                    if torch.is_tensor(v):
                        live[k][name].copy_(v)
            if not saved_opt["state"]:  # fresh optimizer: zero the moments and the step counters the warm-up created
                for stt in live.values():
                    for v in stt.values():
                        if torch.is_tensor(v):
                            v.zero_()

    def step(self):
        """One iteration; no host synchronisation.  self.loss / self.warped hold its loss and warped source."""
        with torch.cuda.device(self.source.device):
            if not self.use_graph:
                for g in self.optimizer.param_groups:
                    for p in g["params"]:
                        if p.grad is None:
                            p.grad = torch.zeros_like(p)
                self._body()
            else:
                if self.graph is None:
                    self._capture()
                self.graph.replay()
        self.iterations += 1

    def run(self, iters: int, read_every: int = 10, loss_tol: float = 1e-4, break_threshold_ratio: float = 1e-4,
            max_break_count: int = 15):
        """The loop of registration.py:210-260 with its early-stop rules evaluated on every read_every-th loss.
        Returns the list of (iteration, loss) pairs that were read."""
        hist, loss_prev, break_counter = [], 1e6, 0
        for it in range(iters):
            self.step()
            if it % read_every == 0 or it == iters - 1:
                val = float(self.loss.item())
                hist.append((it, val))
                if val < loss_tol:
                    break
                if abs(loss_prev - val) < loss_prev * break_threshold_ratio:
                    break_counter += 1
                if break_counter >= max_break_count:
                    break
                loss_prev = val
        return hist
